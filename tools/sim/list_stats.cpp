// list_stats.cpp — DEVELOPMENT TOOL: non-zero coefficients per write-pass thread (half a subsequence),
// to size the per-thread lists of huff_write.
#include "huff_lut.hpp"
#include "huff_threads.hpp"
#include "parser.hpp"
#include <algorithm>
#include <cstdio>
#include <vector>
using namespace jb2;
struct CountMem : HostMem {
    mutable int puts = 0;
    void put_symbol(uint32_t e, uint32_t peek, uint32_t qq) const { put_symbol_generic(*this, e, peek, qq); }
    void put(uint32_t, int) const { ++puts; }
    template <bool WRITE>
    uint32_t step(uint32_t e, uint32_t& pp, uint32_t& qq, uint32_t& ch, uint32_t& tbl, const SlotRec& r,
                  uint32_t& w0, uint32_t& w1, uint32_t& w2, uint32_t& wn, uint32_t& peek, uint32_t end,
                  uint32_t end_fast, uint32_t q_limit) const
    {
        return generic_step<WRITE>(*this, e, pp, qq, ch, tbl, r, w0, w1, w2, wn, peek, end, end_fast, q_limit);
    }
};
int main(int argc, char** argv)
{
    FILE* f = fopen(argv[1], "rb");
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> data(sz); if (fread(data.data(), 1, sz, f) != (size_t)sz) return 1;
    static ImagePlan plan; Logger log;
    if (parse_jpeg(data.data(), data.size(), plan, log) != JPEGGPU_SUCCESS) return 1;
    ScanPlan& sp = plan.scans[0]; ScanGeom& g = sp.geom;
    std::vector<uint32_t> pool; HuffSlow slow[kMaxTables];
    for (int t = 0; t < g.num_tables; ++t) g.table_off[t] = build_lut(sp.tables[t], pool, slow[t]);
    SlotRec recs[kMaxDuPerMcu], recsw[kMaxDuPerMcu]; uint32_t tbls[kMaxDuPerMcu];
    make_slot_recs(g, recs, tbls); make_slot_recs(g, recsw, tbls, true);
    const uint32_t N = g.num_subs;
    std::vector<uint8_t> dst((size_t)N * kSubBytes + 64, 0);
    { const uint8_t* src = data.data() + sp.begin; uint32_t kept = 0, seg = 0, prev = 0;
      for (uint32_t i = 0; i < g.scan_bytes; ++i) { uint32_t b = src[i]; bool aff = prev == 0xff;
        if ((aff ? b == 0 : b != 0xff) && (int)seg < g.num_segments) { const SegDesc& sd = sp.segs[seg];
          dst[(kept - sd.kept_before + sd.sub_off * (uint32_t)kSubBytes) ^ 3u] = aff ? 0xff : (uint8_t)b; ++kept; }
        if (aff && b >= 0xd0 && b <= 0xd7) ++seg; prev = b; } }
    const uint32_t* words = reinterpret_cast<const uint32_t*>(dst.data());
    std::vector<int> counts;
    for (int k = 0; k < g.num_segments; ++k) {
        const SegDesc& seg = sp.segs[k];
        HostMem base{pool.data(), recs, tbls, slow, words + (size_t)seg.sub_off * kSubWords, g.table_off};
        uint16_t st = 0; uint32_t q0 = 0;
        const uint32_t n = subs_in_segment(seg);
        for (uint32_t i = 0; i < n; ++i) {
            Mids mids; const unsigned long long rec = decode_sub(base, seg.nbytes * 8u, i, st, &mids);
            for (uint32_t h = 0; h < 2; ++h) {
                CountMem cm; static_cast<HostMem&>(cm) = base; cm.recs_ = recsw;
                static int16_t sink_c[1 << 20]; (void)sink_c;
                write_part(cm, seg, i, h, rec, h ? mids.v[0] : 0u, q0, 0xffffffffu);
                counts.push_back(cm.puts);
            }
            q0 += entry_n(rec); st = entry_x(rec);
        }
    }
    std::sort(counts.begin(), counts.end());
    double sum = 0; for (int c : counts) sum += c;
    auto pct = [&](double p) { return counts[(size_t)(p * (counts.size() - 1))]; };
    printf("%s: threads %zu mean %.1f p50 %d p90 %d p99 %d p99.9 %d max %d | over 40: %.1f%% 48: %.1f%% 56: %.1f%% 64: %.1f%% 72: %.1f%% 80: %.2f%% 96: %.2f%%\n", argv[1], counts.size(), sum / counts.size(),
           pct(.5), pct(.9), pct(.99), pct(.999), counts.back(),
           100.0 * (counts.end() - std::upper_bound(counts.begin(), counts.end(), 40)) / counts.size(),
           100.0 * (counts.end() - std::upper_bound(counts.begin(), counts.end(), 48)) / counts.size(),
           100.0 * (counts.end() - std::upper_bound(counts.begin(), counts.end(), 56)) / counts.size(),
           100.0 * (counts.end() - std::upper_bound(counts.begin(), counts.end(), 64)) / counts.size(),
           100.0 * (counts.end() - std::upper_bound(counts.begin(), counts.end(), 72)) / counts.size(),
           100.0 * (counts.end() - std::upper_bound(counts.begin(), counts.end(), 80)) / counts.size(),
           100.0 * (counts.end() - std::upper_bound(counts.begin(), counts.end(), 96)) / counts.size());
    return 0;
}
