// emul.cpp — TEST INFRASTRUCTURE.  Runs the product's host+device shared code (parser, LUT
// builder, decode_span and the per-thread bodies of the Huffman kernels in huff_threads.hpp)
// sequentially on the CPU, so the sync/verify/write logic can be checked against the oracle on
// a machine without a GPU.  The kernels' glue that cannot run here (warp scans, shared memory
// staging) is replaced by plain loops that state the same contract.
#include "huff_lut.hpp"
#include "huff_threads.hpp"
#include "parser.hpp"

#include <cstring>
#include <vector>

using namespace jb2;

namespace {
const uint8_t kNat[64] = {JB2_NATURAL_ORDER};
}

extern "C" {

// Parse only: returns status; fills a few numbers for tests.
int emul_parse(const uint8_t* data, size_t size, int* out /* [16] */)
{
    static ImagePlan plan;
    Logger log;
    jpeggpu_status st = parse_jpeg(data, size, plan, log);
    if (st != JPEGGPU_SUCCESS) return (int)st;
    out[0] = plan.width;
    out[1] = plan.height;
    out[2] = plan.num_comp;
    out[3] = plan.num_scans;
    out[4] = plan.restart_interval;
    out[5] = plan.scans[0].geom.num_segments;
    out[6] = (int)plan.scans[0].geom.num_subs;
    out[7] = (int)plan.scans[0].geom.scan_bytes;
    out[8] = (int)plan.scans[0].begin;
    out[9] = (int)plan.scans[0].end;
    return 0;
}

// Full emulated decode of scan `scan_idx` into DU-major coded-order coefficients (DC absolute).
// coef must hold total_mcus * dus_per_mcu * 64 int16.  destuffed_out (optional) receives the
// destuffed buffer in natural byte order (num_subs * kSubBytes bytes).
// stats: [0..2] repairs per pass, [3] num_subs, [4] records still inconsistent before fixup,
//        [5] number of speculate guesses that were wrong.
int emul_decode_scan(const uint8_t* data, size_t size, int scan_idx, int16_t* coef,
                     size_t coef_count, uint8_t* destuffed_out, size_t destuffed_cap, int* stats)
{
    static ImagePlan plan;
    Logger log;
    jpeggpu_status st = parse_jpeg(data, size, plan, log);
    if (st != JPEGGPU_SUCCESS) return (int)st;
    if (scan_idx >= plan.num_scans) return -1;
    ScanPlan& sp = plan.scans[scan_idx];
    ScanGeom& g = sp.geom;
    const size_t need = (size_t)g.total_mcus * g.dus_per_mcu * 64;
    if (coef_count < need) return -2;

    std::vector<uint16_t> pool;
    HuffSlow slow[kMaxTables];
    for (int t = 0; t < g.num_tables; ++t) g.table_off[t] = (uint16_t)build_lut(sp.tables[t], pool, slow[t]);
    g.lut_pool_u16 = (uint32_t)pool.size();
    uint32_t offs[kMaxDuPerMcu], tbls[kMaxDuPerMcu];
    for (int s = 0; s < kMaxDuPerMcu; ++s) {
        const int ss = s < g.dus_per_mcu ? s : 0;
        offs[s] = (uint32_t)g.table_off[g.slot_dc[ss]] | ((uint32_t)g.table_off[g.slot_ac[ss]] << 16);
        tbls[s] = (uint32_t)g.slot_dc[ss] | ((uint32_t)g.slot_ac[ss] << 8);
    }
    const HostMem T{pool.data(), offs, tbls, slow, kNat};
    const SlotTables K = make_slot_tables(T, (uint32_t)g.dus_per_mcu);

    // destuff: the contract of destuff_count + destuff_apply
    const uint32_t N = g.num_subs;
    std::vector<uint8_t> dst((size_t)N * kSubBytes + 64, 0);
    {
        const uint8_t* src = data + sp.begin;
        uint32_t kept = 0, seg = 0, prev = 0;
        for (uint32_t i = 0; i < g.scan_bytes; ++i) {
            const uint32_t b = src[i];
            const bool after_ff = prev == 0xff;
            const bool keep = after_ff ? b == 0 : b != 0xff;
            const bool rst = after_ff && b >= 0xd0 && b <= 0xd7;
            if (keep) {
                if ((int)seg >= g.num_segments) return -3;
                const uint32_t adj = sp.segs[seg].sub_off * (uint32_t)kSubBytes - sp.segs[seg].kept_before;
                const uint32_t d = kept + adj;
                if (d >= dst.size()) return -4;
                dst[d ^ 3u] = after_ff ? 0xff : (uint8_t)b;
                ++kept;
            }
            if (rst) ++seg;
            prev = b;
        }
    }
    if (destuffed_out) {
        const size_t n = (size_t)N * kSubBytes;
        if (destuffed_cap < n) return -5;
        for (size_t d = 0; d < n; ++d) destuffed_out[d] = dst[d ^ 3u];
    }
    const uint32_t* words = reinterpret_cast<const uint32_t*>(dst.data());

    std::vector<unsigned long long> entries(N);
    std::vector<uint32_t> seg_idx(N), qpos(N + 1);
    for (uint32_t j = 0; j < N; ++j) entries[j] = speculate_one(T, K, sp.segs.data(), g.num_segments, words, j, &seg_idx[j]);
    // count wrong guesses
    int wrong = 0;
    for (uint32_t j = 0; j < N; ++j) {
        const SegDesc& sg = sp.segs[seg_idx[j]];
        if (j != sg.sub_off && (uint16_t)(entries[j - 1] >> 16) != (uint16_t)entries[j]) ++wrong;
    }
    stats[5] = wrong;
    uint32_t last = 1;
    for (int pass = 0; pass < 3; ++pass) {
        stats[pass] = 0;
        if (pass > 0 && last == 0) continue;
        // emulate concurrency pessimistically: every thread decides from a snapshot
        std::vector<unsigned long long> snap(entries);
        uint32_t total = 0;
        for (uint32_t j = 0; j < N; ++j) {
            const SegDesc& sg = sp.segs[seg_idx[j]];
            if (j == sg.sub_off) continue;
            if ((uint16_t)(snap[j - 1] >> 16) == (uint16_t)snap[j]) continue;
            total += repair_one(T, K, sp.segs.data(), seg_idx.data(), words, entries.data(), j);
        }
        stats[pass] = (int)total;
        last = total;
    }
    int bad = 0;
    for (uint32_t j = 0; j < N; ++j) {
        const SegDesc& sg = sp.segs[seg_idx[j]];
        if (j != sg.sub_off && (uint16_t)(entries[j - 1] >> 16) != (uint16_t)entries[j]) ++bad;
    }
    stats[3] = (int)N;
    stats[4] = bad;
    if (last != 0)
        for (int k = 0; k < g.num_segments; ++k) fixup_segment(T, K, sp.segs[k], words, entries.data());
    uint32_t acc = 0;
    for (uint32_t j = 0; j < N; ++j) {
        qpos[j] = acc;
        acc += (uint32_t)(entries[j] >> 32);
    }
    qpos[N] = acc;
    memset(coef, 0, need * sizeof(int16_t));
    for (uint32_t j = 0; j < N; ++j)
        write_one(T, K, g, sp.segs.data(), seg_idx.data(), words, entries.data(), qpos.data(), coef, j);
    // DC: the contract of dc_tile_sums + dc_apply
    for (int a = 0; a < g.num_comp; ++a) {
        int pred = 0;
        for (int m = 0; m < g.total_mcus; ++m) {
            if (m % g.restart_interval == 0) pred = 0;
            for (int s = 0; s < g.dus_per_mcu; ++s)
                if (g.slot_comp[s] == a) {
                    int16_t* du = coef + ((size_t)m * g.dus_per_mcu + s) * 64;
                    pred += du[0];
                    du[0] = (int16_t)pred;
                }
        }
    }
    return 0;
}

} // extern "C"
