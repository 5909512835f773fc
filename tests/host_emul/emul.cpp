// emul.cpp — TEST INFRASTRUCTURE.  Runs the product's host+device shared code (parser, LUT
// builder, decode_span and the per-thread bodies of the Huffman kernels in huff_threads.hpp)
// sequentially on the CPU, so the sync/verify/write logic can be checked against the oracle on
// a machine without a GPU.  The kernels' glue that cannot run here (warp scans, shared memory
// staging) is replaced by plain loops that state the same contract.
#include "huff_lut.hpp"
#include "huff_threads.hpp"
#include "parser.hpp"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace jb2;

namespace {
const uint8_t kNat[64] = {JB2_NATURAL_ORDER};
}

struct SinglesOnly : HostMem { // slot records point at the singles halves; single_of() stays there
    JB2_HD uint32_t single_of(uint32_t tbl) const { return tbl; }
    template <bool WRITE>
    JB2_HD uint32_t step(uint32_t e, uint32_t& pp, uint32_t& qq, uint32_t& ch, uint32_t& tbl,
                         const SlotRec& r, uint32_t& w0, uint32_t& w1, uint32_t& w2, uint32_t& wn,
                         uint32_t& peek, uint32_t end, uint32_t end_fast, uint32_t q_limit) const
    {
        return generic_step<WRITE>(*this, e, pp, qq, ch, tbl, r, w0, w1, w2, wn, peek, end, end_fast, q_limit);
    }
};

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
// stats: [0] chunk boundaries healed inside huff_sync, [1] most fix-point rounds of any warp,
//        [2] records rewritten by huff_link's fallback, [3] num_subs, [4] link rounds;
//        in: stats[7] == 12345 corrupts records before huff_link (fallback test hook).
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

    std::vector<uint32_t> pool;
    HuffSlow slow[kMaxTables];
    for (int t = 0; t < g.num_tables; ++t) g.table_off[t] = build_lut(sp.tables[t], pool, slow[t]);
    g.lut_pool_words = (uint32_t)pool.size();
    SlotRec recs[kMaxDuPerMcu], recs_write[kMaxDuPerMcu];
    uint32_t tbls[kMaxDuPerMcu];
    make_slot_recs(g, recs, tbls);
    make_slot_recs(g, recs_write, tbls, true); // the write pass decodes with the singles halves

    // destuff: the contract of destuff_kernel, tile by tile from the host tile directory
    const uint32_t N = g.num_subs;
    std::vector<uint8_t> dst((size_t)N * kSubBytes + 64, 0xa5); // poison: padding must be written
    {
        const uint8_t* src = data + sp.begin;
        const uint32_t tiles = (g.scan_bytes + kDestuffTile - 1) / kDestuffTile;
        if (sp.tiles.size() != tiles) return -6;
        for (uint32_t t = 0; t < tiles; ++t) {
            uint32_t kept = sp.tiles[t].kept_before, seg = sp.tiles[t].segs_before;
            const uint32_t i0 = t * kDestuffTile, i1 = umin(i0 + kDestuffTile, g.scan_bytes);
            uint32_t prev = i0 ? src[i0 - 1] : 0u;
            for (uint32_t i = i0; i < i1; ++i) {
                const uint32_t b = src[i];
                const bool after_ff = prev == 0xff;
                const bool keep = after_ff ? b == 0 : b != 0xff;
                const bool rst = after_ff && b >= 0xd0 && b <= 0xd7;
                if (keep) {
                    if ((int)seg >= g.num_segments) return -3;
                    const SegDesc& sd = sp.segs[seg];
                    const uint32_t r = kept - sd.kept_before;
                    const uint32_t d = r + sd.sub_off * (uint32_t)kSubBytes;
                    if (d >= dst.size()) return -4;
                    dst[d ^ 3u] = after_ff ? 0xff : (uint8_t)b;
                    if (r + 1 == sd.nbytes) {
                        const uint32_t pad_end = (d + kSubBytes) / kSubBytes * kSubBytes;
                        for (uint32_t z = d + 1; z < pad_end; ++z) dst[z ^ 3u] = 0;
                    }
                    ++kept;
                }
                if (rst) ++seg;
                prev = b;
            }
        }
    }
    if (destuffed_out) {
        const size_t n = (size_t)N * kSubBytes;
        if (destuffed_cap < n) return -5;
        for (size_t d = 0; d < n; ++d) destuffed_out[d] = dst[d ^ 3u];
    }
    const uint32_t* words = reinterpret_cast<const uint32_t*>(dst.data());
    long long_codes = 0;
    auto mem_of = [&](const SegDesc& seg) {
        HostMem m{pool.data(), recs, tbls, slow, words + (size_t)seg.sub_off * kSubWords, g.table_off};
        m.long_codes_ = &long_codes;
        return m;
    };

    // huff_sync: one "warp" per chunk, lanes emulated by arrays; chunks in launch order so the
    // left mailbox is always there (one valid schedule of the kernel)
    std::vector<unsigned long long> entries(kRecWords * (size_t)N); // kRecWords words per record
    auto rec_lo = [&](uint32_t j) -> unsigned long long& { return entries[kRecWords * (size_t)j]; };
    std::vector<uint32_t> seg_idx(N), qpos(N + 1);
    const uint32_t chunks = (N + kChunkSubs - 1) / kChunkSubs;
    std::vector<uint32_t> mailbox(chunks, 0);
    int relinked = 0, rounds_max = 0, rounds_sum = 0;
    long diag[3] = {0, 0, 0};
    int passes_max = 0;
    for (uint32_t chunk = 0; chunk < chunks; ++chunk) {
        const uint32_t chunk_first = chunk * kChunkSubs;
        bool active[32], head[32];
        uint32_t k[32], i[32], seg_bits[32];
        uint16_t entry[32];
        unsigned long long rec[32];
        Mids mid[32];
        uint32_t memo_entry[32]; // the one-entry memo of sync_chunk (kernels_huffman.cu)
        unsigned long long memo_rec[32];
        Mids memo_mid[32];
        for (int l = 0; l < 32; ++l) {
            const long long sub = (long long)chunk_first - kOverlapSubs + l;
            active[l] = sub >= 0 && sub < (long long)N;
            head[l] = false;
            entry[l] = 0;
            rec[l] = 0;
            mid[l] = Mids{};
            memo_entry[l] = 0xffffffffu;
            memo_rec[l] = 0;
            memo_mid[l] = Mids{};
            if (!active[l]) continue;
            k[l] = find_segment(sp.segs.data(), g.num_segments, (uint32_t)sub);
            const SegDesc& seg = sp.segs[k[l]];
            i[l] = (uint32_t)sub - seg.sub_off;
            head[l] = i[l] == 0;
            seg_bits[l] = seg.nbytes * 8u;
            rec[l] = decode_sub(mem_of(seg), seg_bits[l], i[l], 0, &mid[l]);
        }
        int chunk_passes = 1;
        auto adopt = [&](int l, uint16_t want) { // returns true when a decode was needed
            const uint32_t old_entry = entry[l];
            const unsigned long long old_rec = rec[l];
            const Mids old_mid = mid[l];
            entry[l] = want;
            const bool hit = memo_entry[l] == (uint32_t)want;
            if (hit) {
                rec[l] = memo_rec[l];
                mid[l] = memo_mid[l];
            } else {
                rec[l] = decode_sub(mem_of(sp.segs[k[l]]), seg_bits[l], i[l], entry[l], &mid[l]);
            }
            memo_entry[l] = old_entry;
            memo_rec[l] = old_rec;
            memo_mid[l] = old_mid;
            return !hit;
        };
        auto iterate = [&](int frozen) {
            int rounds = 0;
            for (;;) {
                uint16_t want[32];
                bool any = false;
                for (int l = 0; l < 32; ++l) {
                    const uint16_t px = l ? entry_x(rec[l - 1]) : 0;
                    want[l] = (l <= frozen || head[l] || !active[l]) ? entry[l] : px;
                    any |= want[l] != entry[l];
                }
                if (!any) break;
                bool decoded = false;
                for (int l = 0; l < 32; ++l)
                    if (want[l] != entry[l]) {
                        if (rounds >= 4 && getenv("EMUL_DIAG")) {
                            const uint16_t d = want[l] ^ entry[l];
                            ++diag[(d & 0x1f) ? 0 : (d & 0x7e00) ? 1 : 2]; // o differs / z differs / c only
                        }
                        decoded |= adopt(l, want[l]);
                    }
                if (decoded) ++rounds; // passes in which some lane had to decode
            }
            if (rounds >= 100 && getenv("EMUL_DIAG")) {
                fprintf(stderr, "chunk %u rounds %d frozen %d n:", chunk, rounds, frozen);
                for (int l = 0; l < 32; l += 3) fprintf(stderr, " %u", entry_n(rec[l]));
                fprintf(stderr, "\n");
            }
            if (rounds > rounds_max) rounds_max = rounds;
            rounds_sum += rounds;
            chunk_passes += rounds;
        };
        iterate(0);
        const uint32_t last_sub = umin(chunk_first + kChunkSubs, N) - 1;
        const int last_lane = (int)(last_sub - chunk_first) + kOverlapSubs;
        mailbox[chunk] = 0x80000000u | entry_x(rec[last_lane]);
        if (chunk > 0 && !head[kOverlapSubs]) {
            const uint32_t left = mailbox[chunk - 1];
            if ((uint16_t)(left & 0xffffu) != entry[kOverlapSubs]) {
                adopt(kOverlapSubs, (uint16_t)(left & 0xffffu));
                ++relinked;
                iterate(kOverlapSubs);
                mailbox[chunk] = 0x80000000u | entry_x(rec[last_lane]);
            }
        }
        if (chunk_passes > passes_max) passes_max = chunk_passes;
        for (int l = kOverlapSubs; l < 32; ++l)
            if (active[l]) {
                const uint32_t sub = chunk_first - kOverlapSubs + l;
                rec_lo(sub) = pack_entry((uint16_t)(entry[l] | (head[l] ? kStateHead : 0)),
                                         entry_x(rec[l]), entry_n(rec[l]));
                for (int w = 0; 2 * w < kWriteParts - 1; ++w)
                    entries[kRecWords * (size_t)sub + 1 + w] = mids_word(mid[l], w);
                seg_idx[sub] = k[l];
            }
    }
    if (getenv("EMUL_DIAG")) fprintf(stderr, "late changes: o %ld, z(not o) %ld, c only %ld\n", diag[0], diag[1], diag[2]);
    stats[0] = relinked;
    stats[1] = rounds_max;
    stats[5] = rounds_sum;
    stats[6] = (int)chunks;
    if (getenv("EMUL_DIAG")) fprintf(stderr, "S=%d R=%d chunks %u relinked %d passes_max %d rounds_avg %.2f\n", kSubBytes, kOverlapSubs, chunks, relinked, passes_max, (double)rounds_sum / chunks);
    stats[3] = (int)N;

    // huff_link: verify chunk boundaries, sequential fallback, segmented exclusive scan
    if (stats[7] == 12345) // test hook: corrupt some records to exercise the fallback
        for (uint32_t j = 3; j < N; j += 7)
            if (!(entry_g(rec_lo(j)) & kStateHead)) rec_lo(j) = pack_entry(entry_g(rec_lo(j)) ^ 0x21, entry_x(rec_lo(j)) ^ 0x40, 5);
    uint32_t stride = stats[7] == 12345 ? 1 : kChunkSubs;
    int slow_fix = 0, link_rounds = 0;
    for (;;) {
        std::vector<unsigned long long> snap(entries); // every thread decides from a snapshot
        uint32_t bad = 0;
        for (uint32_t j = stride; j < N; j += stride)
            if (!boundary_ok(snap[kRecWords * (size_t)(j - 1)], snap[kRecWords * (size_t)j])) {
                ++bad;
                const SegDesc& seg = sp.segs[seg_idx[j]];
                slow_fix += (int)relink_from(mem_of(seg), seg, entries.data(), j);
            }
        if (!bad) break;
        stride = 1;
        ++link_rounds;
    }
    stats[2] = slow_fix;
    stats[4] = link_rounds;
    for (uint32_t j = 1; j < N; ++j)
        if (!boundary_ok(rec_lo(j - 1), rec_lo(j))) return -7; // inconsistent chain
    uint32_t acc = 0;
    for (uint32_t j = 0; j < N; ++j) {
        if (entry_g(rec_lo(j)) & kStateHead) acc = 0;
        qpos[j] = acc;
        acc += entry_n(rec_lo(j));
    }

    // huff_write
    long_codes = 0;
    std::vector<int16_t> zz(need, 0); // DU-major, zig-zag order inside a DU (the device layout)
    std::vector<int16_t> dc((size_t)g.total_mcus * g.dus_per_mcu, 0);
    for (uint32_t j = 0; j < N; ++j) {
        const SegDesc& seg = sp.segs[seg_idx[j]];
        uint32_t first_du, q_limit;
        if (!segment_extent(g, seg_idx[j], first_du, q_limit)) continue;
        HostMem m = mem_of(seg);
        m.recs_ = recs_write;
        m.coef_ = zz.data() + (size_t)first_du * 64;
        m.dc_ = dc.data() + first_du;
        for (uint32_t h = 0; h < (uint32_t)kWriteParts; ++h) // the threads of a subsequence
            write_part(m, seg, j - seg.sub_off, h, rec_lo(j),
                       h ? mid_of(&entries[kRecWords * (size_t)j + 1], (int)h - 1) : 0u, qpos[j], q_limit);
    }
    if (getenv("EMUL_DIAG")) fprintf(stderr, "lut bits %d: long codes in the write pass %ld\n", kLutBits, long_codes);
    // DC: the contract of dc_tile_sums + dc_apply
    for (int a = 0; a < g.num_comp; ++a) {
        int pred = 0;
        for (int m = 0; m < g.total_mcus; ++m) {
            if (m % g.restart_interval == 0) pred = 0;
            for (int s = 0; s < g.dus_per_mcu; ++s)
                if (g.slot_comp[s] == a) {
                    const size_t du = (size_t)m * g.dus_per_mcu + s;
                    pred += dc[du];
                    zz[du * 64] = (int16_t)pred;
                }
        }
    }
    // the oracle's stream layout is raster order inside a DU: un-zig-zag (IDCT / transpose do
    // this on the device)
    for (size_t du = 0; du < need / 64; ++du)
        for (int z = 0; z < 64; ++z) coef[du * 64 + kNat[z]] = zz[du * 64 + z];
    return 0;
}

// Pair-step equivalence: the state-only pass consumes two symbols per probe inside its fast zone
// (huff_core.hpp decode_span<false>); the result must equal single stepping for ANY entry state,
// also the speculative ones the fix-point feeds it.  Decodes every subsequence of scan 0 from the
// cold state and from `tries` pseudo-random entry states with both and returns the number of
// differing records (0 expected), or a negative parse error.
int emul_pair_equivalence(const uint8_t* data, size_t size, int tries, long* pairs_possible)
{
    static ImagePlan plan;
    Logger log;
    if (parse_jpeg(data, size, plan, log) != JPEGGPU_SUCCESS) return -1;
    ScanPlan& sp = plan.scans[0];
    ScanGeom& g = sp.geom;
    std::vector<uint32_t> pool;
    HuffSlow slow[kMaxTables];
    for (int t = 0; t < g.num_tables; ++t) g.table_off[t] = build_lut(sp.tables[t], pool, slow[t]);
    long pairs = 0;
    for (uint32_t e : pool) pairs += e >= kEntryMarkerLimit && entry_skip2(e) != 0;
    if (pairs_possible) *pairs_possible = pairs;
    SlotRec recs[kMaxDuPerMcu], recs_single[kMaxDuPerMcu];
    uint32_t tbls[kMaxDuPerMcu];
    make_slot_recs(g, recs, tbls);
    make_slot_recs(g, recs_single, tbls, true);
    const uint32_t N = g.num_subs;
    std::vector<uint8_t> dst((size_t)N * kSubBytes + 64, 0);
    {
        const uint8_t* src = data + sp.begin;
        uint32_t kept = 0, seg = 0, prev = 0;
        for (uint32_t i = 0; i < g.scan_bytes; ++i) {
            const uint32_t b = src[i];
            const bool after_ff = prev == 0xff;
            if ((after_ff ? b == 0 : b != 0xff) && (int)seg < g.num_segments) {
                const SegDesc& sd = sp.segs[seg];
                dst[(kept - sd.kept_before + sd.sub_off * (uint32_t)kSubBytes) ^ 3u] = after_ff ? 0xff : (uint8_t)b;
                ++kept;
            }
            if (after_ff && b >= 0xd0 && b <= 0xd7) ++seg;
            prev = b;
        }
    }
    const uint32_t* words = reinterpret_cast<const uint32_t*>(dst.data());
    int bad = 0;
    uint32_t rng = 12345u;
    for (int k = 0; k < g.num_segments; ++k) {
        const SegDesc& seg = sp.segs[k];
        HostMem pm{pool.data(), recs, tbls, slow, words + (size_t)seg.sub_off * kSubWords, g.table_off};
        SinglesOnly sm;
        static_cast<HostMem&>(sm) = pm;
        sm.recs_ = recs_single;
        const uint32_t n = subs_in_segment(seg);
        for (uint32_t i = 0; i < n; ++i)
            for (int t = 0; t <= tries; ++t) {
                rng = rng * 1664525u + 1013904223u;
                const uint16_t gs = t == 0 ? 0 : pack_state((rng >> 8) % 32u, (rng >> 13) % (uint32_t)g.dus_per_mcu, (rng >> 20) % 64u);
                Mids ma, mb;
                const unsigned long long a = decode_sub(pm, seg.nbytes * 8u, i, gs, &ma);
                const unsigned long long b = decode_sub(sm, seg.nbytes * 8u, i, gs, &mb);
                bool same = a == b;
                for (int w = 0; w < kWriteParts - 1; ++w) same &= ma.v[w] == mb.v[w];
                bad += same ? 0 : 1;
            }
    }
    return bad;
}

// Walk equivalence: eager parse vs deferred parse + finish, sequential (threads = 0) or with the
// helper pool (threads = 1).  Returns 0 when both give the same status and - on success - the
// same geometry, segment directory and tile directory; otherwise a code naming the difference.
// *used_pool = 1 when the scan was large enough for the pool to be kicked.
int emul_walk_compare(const uint8_t* data, size_t size, int threads, int* used_pool, int* status)
{
    static ImagePlan eager, lazy;
    static WalkPool* pool = walk_pool_create();
    Logger log;
    const jpeggpu_status se = parse_jpeg(data, size, eager, log, false);
    jpeggpu_status sl = parse_jpeg(data, size, lazy, log, true);
    *used_pool = 0;
    if (sl == JPEGGPU_SUCCESS && lazy.walk_pending) {
        if (threads && size - lazy.scans[0].begin >= kWalkParallelMin) {
            walk_pool_kick(pool, lazy);
            *used_pool = 1;
        }
        const uint64_t before = walk_pool_stitched(pool);
        sl = finish_deferred_walk(lazy, log, threads ? pool : nullptr);
        if (*used_pool && walk_pool_stitched(pool) != before) *used_pool = 2; // parallel result kept
    }
    *status = (int)sl;
    if (se != sl) return 1;
    if (se != JPEGGPU_SUCCESS) return 0;
    if (eager.num_scans != lazy.num_scans) return 2;
    for (int s = 0; s < eager.num_scans; ++s) {
        const ScanPlan &a = eager.scans[s], &b = lazy.scans[s];
        if (a.begin != b.begin || a.end != b.end) return 3;
        if (memcmp(&a.geom, &b.geom, sizeof(ScanGeom)) != 0) return 4;
        if (a.segs.size() != b.segs.size() ||
            memcmp(a.segs.data(), b.segs.data(), a.segs.size() * sizeof(SegDesc)) != 0)
            return 5;
        if (a.tiles.size() != b.tiles.size() ||
            memcmp(a.tiles.data(), b.tiles.data(), a.tiles.size() * sizeof(TileDesc)) != 0)
            return 6;
    }
    return 0;
}

} // extern "C"
