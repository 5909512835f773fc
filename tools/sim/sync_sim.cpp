// sync_sim.cpp — DEVELOPMENT TOOL (not product, not test): replays the huff_sync warp fix-point on
// the CPU with a cost model (symbol-loop steps on the critical path) so that variants of the
// iteration can be compared without a GPU.  Uses the product's parser, LUT builder and
// decode_span.  Build: see tools/sim/Makefile.
#include "huff_lut.hpp"
#include "huff_threads.hpp"
#include "parser.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

using namespace jb2;

static long g_steps = 0, g_long = 0;
struct SimMem {
    const uint32_t* lut_;
    const SlotRec* recs_;
    const uint32_t* tbls_;
    const HuffSlow* slow_;
    const uint32_t* words_;
    const uint32_t* table_off_;
    static constexpr uint32_t kWordStep = 1;
    uint32_t single_of(uint32_t tbl) const { return tbl + (uint32_t)kLut1Size; }
    uint32_t cont_table(uint32_t e) const { return entry_cont_vec16(e) * 4u; }
    uint32_t lut(uint32_t tbl, uint32_t i) const
    {
        if (i < (uint32_t)kLut1Size) ++g_steps;
        return lut_[(uint32_t)(tbl + i)];
    }
    uint32_t slot_handle(uint32_t c) const { return c; }
    SlotRec rec(uint32_t h) const { return recs_[h]; }
    uint32_t slot_tbls(uint32_t c) const { return tbls_[c]; }
    const HuffSlow& slow(uint32_t s) const { return slow_[s]; }
    uint32_t word_handle(uint32_t i) const { return i; }
    uint32_t word(uint32_t h) const { return words_[h]; }
    uint32_t probe(uint32_t tbl, uint32_t peek) const { return lut(tbl, peek >> (32 - kLutBits)); }
    uint32_t slide_peek(uint32_t& w0, uint32_t& w1, uint32_t& w2, uint32_t& wn, uint32_t pp, uint32_t pn) const
    {
        slide_window_generic(*this, w0, w1, w2, wn, pp, pn);
        return funnel_left_wrap(w0, w1, pn);
    }
    template <bool WRITE>
    uint32_t step(uint32_t e, uint32_t& pp, uint32_t& qq, uint32_t& ch, uint32_t& tbl, const SlotRec& r,
                  uint32_t& w0, uint32_t& w1, uint32_t& w2, uint32_t& wn, uint32_t& peek, uint32_t end,
                  uint32_t end_fast, uint32_t q_limit) const
    {
        return generic_step<WRITE>(*this, e, pp, qq, ch, tbl, r, w0, w1, w2, wn, peek, end, end_fast, q_limit);
    }
    void put_symbol(uint32_t e, uint32_t peek, uint32_t qq) const { put_symbol_generic(*this, e, peek, qq); }
    void put(uint32_t, int) const {}
    void note_long_code() const { ++g_long; }
};

struct Scan {
    ScanGeom g;
    std::vector<SegDesc> segs;
    std::vector<uint32_t> pool;
    HuffSlow slow[kMaxTables];
    SlotRec recs[kMaxDuPerMcu];
    uint32_t tbls[kMaxDuPerMcu];
    std::vector<uint8_t> dst;
    std::vector<uint32_t> seg_of; // per sub
    uint32_t N;
};

static int CK = 4; // checkpoints per subsequence (the last one is the exit)

struct Eval {              // result of decoding one subsequence from one entry state
    uint16_t ck[8];        // packed state at every checkpoint (ck[CK-1] = exit state)
    uint16_t steps[8];     // cumulative steps up to the checkpoint
};

struct Sim {
    Scan s;
    std::unordered_map<uint64_t, Eval> memo;
    const Eval& eval(uint32_t j, uint16_t gs)
    {
        const uint64_t key = ((uint64_t)j << 16) | gs;
        auto it = memo.find(key);
        if (it != memo.end()) return it->second;
        const SegDesc& seg = s.segs[s.seg_of[j]];
        SimMem M{s.pool.data(), s.recs, s.tbls, s.slow,
                 reinterpret_cast<const uint32_t*>(s.dst.data()) + (size_t)seg.sub_off * kSubWords, s.g.table_off};
        const uint32_t i = j - seg.sub_off, seg_bits = seg.nbytes * 8u;
        const uint32_t start = i * (uint32_t)kSubBits;
        uint32_t p = start + (gs & 31u), c = (gs >> 5) & 15u, q = (gs >> 9) & 63u;
        Eval e;
        const long s0 = g_steps;
        for (int k = 1; k <= CK; ++k) {
            const uint32_t edge = start + (uint32_t)(k * (kSubBits / CK));
            decode_span<false>(M, p, umin(edge, seg_bits), q, c, 0u);
            e.ck[k - 1] = pack_state(p > edge ? p - edge : 0u, c, q & 63u);
            e.steps[k - 1] = (uint16_t)(g_steps - s0 + 2 * k); // + span set-up
        }
        return memo.emplace(key, e).first->second;
    }
};

static bool load(const char* path, Scan& sc)
{
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    static std::vector<uint8_t> data;
    data.resize(sz);
    if (fread(data.data(), 1, sz, f) != (size_t)sz) return false;
    fclose(f);
    static ImagePlan plan;
    Logger log;
    if (parse_jpeg(data.data(), data.size(), plan, log) != JPEGGPU_SUCCESS) return false;
    ScanPlan& sp = plan.scans[0];
    sc.g = sp.geom;
    sc.segs = sp.segs;
    for (int t = 0; t < sc.g.num_tables; ++t) sc.g.table_off[t] = build_lut(sp.tables[t], sc.pool, sc.slow[t]);
    make_slot_recs(sc.g, sc.recs, sc.tbls);
    sc.N = sc.g.num_subs;
    sc.dst.assign((size_t)sc.N * kSubBytes + 64, 0);
    const uint8_t* src = data.data() + sp.begin;
    uint32_t kept = 0, seg = 0, prev = 0;
    for (uint32_t i = 0; i < sc.g.scan_bytes; ++i) {
        const uint32_t b = src[i];
        const bool after_ff = prev == 0xff;
        const bool keep = after_ff ? b == 0 : b != 0xff;
        const bool rst = after_ff && b >= 0xd0 && b <= 0xd7;
        if (keep && (int)seg < sc.g.num_segments) {
            const SegDesc& sd = sc.segs[seg];
            const uint32_t d = kept - sd.kept_before + sd.sub_off * (uint32_t)kSubBytes;
            sc.dst[d ^ 3u] = after_ff ? 0xff : (uint8_t)b;
            ++kept;
        }
        if (rst) ++seg;
        prev = b;
    }
    sc.seg_of.resize(sc.N);
    for (int k = 0; k < sc.g.num_segments; ++k) {
        const uint32_t n = subs_in_segment(sc.segs[k]);
        for (uint32_t i = 0; i < n; ++i) sc.seg_of[sc.segs[k].sub_off + i] = (uint32_t)k;
    }
    return true;
}

struct Opt {
    bool early = false;  // stop a re-decode at the first checkpoint where it meets the previous trajectory
    bool wait = false;   // do not adopt the exit of a lane that is itself changing this pass
    int memo = 0;        // remembered (entry -> result) pairs per lane beyond the current one
    int hyp = 0;         // 1: pass 0 decodes every MCU-slot hypothesis (extra lanes), picks by left neighbour
    int overlap = kOverlapSubs;
    int chunk = kChunkSubs;
};

struct Stats {
    long chunks = 0, passes_sum = 0, cost_sum = 0, decodes = 0, relink = 0;
    int passes_max = 0;
    long cost_max = 0;
    std::vector<int> pass_hist = std::vector<int>(40, 0);
};

int main(int argc, char** argv)
{
    if (argc < 2) return 1;
    Sim sim;
    if (!load(argv[1], sim.s)) {
        fprintf(stderr, "cannot load %s\n", argv[1]);
        return 1;
    }
    Opt o;
    for (int a = 2; a < argc; ++a) {
        const std::string s = argv[a];
        if (s == "early") o.early = true;
        else if (s == "wait") o.wait = true;
        else if (s.rfind("memo=", 0) == 0) o.memo = atoi(s.c_str() + 5);
        else if (s.rfind("hyp=", 0) == 0) o.hyp = atoi(s.c_str() + 4);
        else if (s.rfind("ck=", 0) == 0) CK = atoi(s.c_str() + 3);
        else if (s.rfind("overlap=", 0) == 0) o.overlap = atoi(s.c_str() + 8);
    }
    o.chunk = 32 - o.overlap;
    Scan& sc = sim.s;
    const uint32_t N = sc.N;
    // truth: sequential decode of every segment
    std::vector<uint16_t> truth(N + 1, 0);
    for (int k = 0; k < sc.g.num_segments; ++k) {
        const uint32_t n = subs_in_segment(sc.segs[k]);
        uint16_t st = 0;
        for (uint32_t i = 0; i < n; ++i) {
            truth[sc.segs[k].sub_off + i] = st;
            st = sim.eval(sc.segs[k].sub_off + i, st).ck[CK - 1];
        }
    }
    // cold-start statistics
    {
        long synced1 = 0, conly = 0, tot = 0;
        for (uint32_t j = 0; j + 1 < N; ++j) {
            if (sc.seg_of[j] != sc.seg_of[j + 1]) continue;
            ++tot;
            const uint16_t x = sim.eval(j, 0).ck[CK - 1];
            if (x == truth[j + 1]) ++synced1;
            else if (((x ^ truth[j + 1]) & ~(15u << 5)) == 0) ++conly;
        }
        long okc = 0, okcz = 0, okcls = 0;
        for (uint32_t j = 0; j + 1 < N; ++j) {
            if (sc.seg_of[j] != sc.seg_of[j + 1]) continue;
            const uint16_t t = truth[j];
            const uint32_t tc = (t >> 5) & 15u, tz = (t >> 9) & 63u;
            // true c of the first block that STARTS in the sub: if z != 0 the current block ends first
            if (sim.eval(j, pack_state(0, tc, 0)).ck[CK - 1] == truth[j + 1]) ++okc;
            if (sim.eval(j, pack_state(0, tc, tz)).ck[CK - 1] == truth[j + 1]) ++okcz;
            const uint32_t nc = tz ? (tc + 1) % sc.g.dus_per_mcu : tc;
            if (sim.eval(j, pack_state(0, nc, 0)).ck[CK - 1] == truth[j + 1]) ++okcls;
        }
        {
            long plur_ok = 0, plur2 = 0, plur2_ok = 0, any_ok = 0, votes_sum = 0;
            const int K = sc.g.dus_per_mcu;
            for (uint32_t j = 0; j + 1 < N; ++j) {
                if (sc.seg_of[j] != sc.seg_of[j + 1]) continue;
                uint16_t ex[16];
                for (int h = 0; h < K; ++h) ex[h] = sim.eval(j, pack_state(0, h, 0)).ck[CK - 1];
                int best = 0, bestv = 0, tv = 0;
                for (int h = 0; h < K; ++h) {
                    int v = 0;
                    for (int h2 = 0; h2 < K; ++h2) v += ex[h2] == ex[h];
                    if (v > bestv) bestv = v, best = h;
                    if (ex[h] == truth[j + 1]) tv = v;
                }
                if (getenv("DUMP") && j % 997 == 5) {
                    fprintf(stderr, "sub %u truth entry (o%u c%u z%u) exit (o%u c%u z%u):", j, truth[j] & 31, (truth[j] >> 5) & 15, truth[j] >> 9, truth[j+1] & 31, (truth[j+1] >> 5) & 15, truth[j+1] >> 9);
                    for (int h = 0; h < K; ++h) fprintf(stderr, "  h%d->(o%u c%u z%u)", h, ex[h] & 31, (ex[h] >> 5) & 15, ex[h] >> 9);
                    fprintf(stderr, "\n");
                }
                votes_sum += tv;
                any_ok += tv > 0;
                plur_ok += ex[best] == truth[j + 1];
                if (bestv >= 2) {
                    ++plur2;
                    plur2_ok += ex[best] == truth[j + 1];
                }
            }
            fprintf(stderr, "K hypotheses: truth among exits %.1f%%, plurality==truth %.1f%%, plurality has >=2 votes %.1f%% (of which right %.2f%%), mean votes for truth %.2f\n",
                    100.0 * any_ok / tot, 100.0 * plur_ok / tot, 100.0 * plur2 / tot, 100.0 * plur2_ok / std::max(1l, plur2), (double)votes_sum / tot);
        }
        fprintf(stderr, "given true c: %.1f%%; true c,z: %.1f%%; c of next block start: %.1f%%\n", 100.0 * okc / tot,
                100.0 * okcz / tot, 100.0 * okcls / tot);
        fprintf(stderr, "N=%u segs=%d K=%d: cold start exact after own sub %.1f%%, wrong only in c %.1f%%\n", N,
                sc.g.num_segments, sc.g.dus_per_mcu, 100.0 * synced1 / tot, 100.0 * conly / tot);
    }
    if (getenv("IDEAL")) {
        // oracle bound: earliest pass at which lane i's Jacobi exit is the truth
        const uint32_t chunks2 = (N + o.chunk - 1) / o.chunk;
        int worst = 0;
        double sum = 0;
        std::vector<int> hist(40, 0);
        for (uint32_t chunk = 0; chunk < chunks2; ++chunk) {
            const long long first = (long long)chunk * o.chunk - o.overlap;
            int tmax = 0;
            for (int l = 0; l < 32; ++l) {
                const long long sub = first + l;
                if (sub < 0 || sub + 1 >= (long long)N) continue;
                if (sc.seg_of[sub] != sc.seg_of[sub + 1]) continue;
                // decoder cold-started k subs before `sub` (not before lane 0 / the segment head)
                int t = -1;
                for (int k = 0; k <= l; ++k) {
                    const long long j0 = sub - k;
                    if (j0 < 0 || sc.seg_of[j0] != sc.seg_of[sub]) { t = k; break; } // head: truth known
                    uint16_t stt = 0;
                    for (long long j = j0; j <= sub; ++j) stt = sim.eval((uint32_t)j, stt).ck[CK - 1];
                    if (stt == truth[sub + 1]) { t = k; break; }
                }
                if (t < 0) t = l + 1;
                tmax = std::max(tmax, t);
            }
            worst = std::max(worst, tmax + 2);
            sum += tmax + 2;
            ++hist[std::min(39, tmax + 2)];
        }
        printf("oracle-selection bound: passes avg %.2f max %d\n", sum / chunks2, worst);
    }
    Stats st;
    const uint32_t chunks = (N + o.chunk - 1) / o.chunk;
    for (uint32_t chunk = 0; chunk < chunks; ++chunk) {
        const long long first = (long long)chunk * o.chunk - o.overlap;
        struct Lane {
            bool active = false, head = false;
            uint32_t j = 0;
            uint16_t entry = 0;
            Eval cur{};
            std::vector<std::pair<uint16_t, Eval>> memo;
        } L[32];
        long cost = 0;
        int passes = 0;
        long pass_cost = 0;
        long span_max[8] = {0};
        auto full_decode = [&](Lane& l, uint16_t g, bool first_time) {
            const Eval& e = sim.eval(l.j, g);
            long steps = e.steps[CK - 1];
            if (!first_time && o.early) {
                for (int k = 0; k < CK; ++k)
                    if (e.ck[k] == l.cur.ck[k]) {
                        steps = e.steps[k];
                        break;
                    }
            }
            if (o.memo) {
                l.memo.insert(l.memo.begin(), {l.entry, l.cur});
                if ((int)l.memo.size() > o.memo) l.memo.pop_back();
            }
            l.entry = g;
            l.cur = e;
            ++st.decodes;
            pass_cost = std::max(pass_cost, steps);
            for (int k = 0; k < CK; ++k)
                span_max[k] = std::max<long>(span_max[k], e.steps[k] - (k ? e.steps[k - 1] : 0));
        };
        auto adopt = [&](Lane& l, uint16_t g) { // returns true if a decode was needed
            for (size_t m = 0; m < l.memo.size(); ++m)
                if (l.memo[m].first == g) {
                    std::swap(l.entry, l.memo[m].first);
                    std::swap(l.cur, l.memo[m].second);
                    return false;
                }
            full_decode(l, g, false);
            return true;
        };
        for (int l = 0; l < 32; ++l) {
            const long long sub = first + l;
            if (sub < 0 || sub >= (long long)N) continue;
            L[l].active = true;
            L[l].j = (uint32_t)sub;
            L[l].head = sc.segs[sc.seg_of[sub]].sub_off == (uint32_t)sub;
        }
        // pass 0
        pass_cost = 0;
        if (o.hyp) {
            // every lane decodes all K hypotheses (c = 0..K-1, o = 0, z = 0) on extra lanes; it then
            // keeps the one whose c matches the exit c of the left lane's c=... (chain from lane 0)
            const int K = sc.g.dus_per_mcu;
            long worst = 0;
            for (int l = 0; l < 32; ++l) {
                if (!L[l].active) continue;
                for (int h = 0; h < K; ++h) worst = std::max<long>(worst, sim.eval(L[l].j, pack_state(0, h, 0)).steps[CK - 1]);
            }
            pass_cost = worst;
            // selection walk: lane 0 takes c=0 (cold), lane l takes hypothesis = c of left's exit
            for (int l = 0; l < 32; ++l) {
                if (!L[l].active) continue;
                uint16_t g = 0;
                if (l > 0 && L[l - 1].active && !L[l].head) g = pack_state(0, (L[l - 1].cur.ck[CK - 1] >> 5) & 15u, 0);
                L[l].entry = g;
                L[l].cur = sim.eval(L[l].j, g);
                st.decodes += K;
            }
        } else {
            for (int l = 0; l < 32; ++l)
                if (L[l].active) full_decode(L[l], 0, true);
        }
        auto close_pass = [&]() {
            long simt = 0;
            for (int k = 0; k < CK; ++k) simt += span_max[k], span_max[k] = 0;
            if (getenv("SIMT")) pass_cost = simt;
            cost += pass_cost + 3;
        };
        close_pass();
        ++passes;
        auto iterate = [&](int frozen) {
            for (;;) {
                uint16_t want[32];
                bool changed[32];
                bool any = false;
                for (int l = 0; l < 32; ++l) {
                    const uint16_t px = l && L[l - 1].active ? L[l - 1].cur.ck[CK - 1] : 0;
                    want[l] = (l <= frozen || L[l].head || !L[l].active) ? L[l].entry : px;
                    changed[l] = want[l] != L[l].entry;
                }
                if (o.wait) {
                    // a lane whose left neighbour is itself changing keeps its entry, unless it still
                    // runs on the cold state (pass 1)
                    bool ch2[32];
                    for (int l = 0; l < 32; ++l)
                        ch2[l] = changed[l] && !(l > 0 && changed[l - 1] && L[l].entry != 0);
                    memcpy(changed, ch2, sizeof(ch2));
                }
                for (int l = 0; l < 32; ++l) any |= changed[l];
                if (!any) break;
                pass_cost = 0;
                bool decoded = false;
                for (int l = 0; l < 32; ++l)
                    if (changed[l]) decoded |= adopt(L[l], want[l]);
                close_pass();
                if (decoded) ++passes;
            }
        };
        iterate(0);
        // mailbox: lane `overlap` must have entered with the truth
        if (L[o.overlap].active && !L[o.overlap].head && L[o.overlap].entry != truth[L[o.overlap].j]) {
            ++st.relink;
            pass_cost = 0;
            adopt(L[o.overlap], truth[L[o.overlap].j]);
            close_pass();
            ++passes;
            iterate(o.overlap);
        }
        for (int l = o.overlap; l < 32; ++l)
            if (L[l].active && L[l].entry != truth[L[l].j]) {
                fprintf(stderr, "chunk %u lane %d: wrong final state\n", chunk, l);
                return 2;
            }
        ++st.chunks;
        st.passes_sum += passes;
        st.cost_sum += cost;
        st.passes_max = std::max(st.passes_max, passes);
        st.cost_max = std::max(st.cost_max, cost);
        ++st.pass_hist[std::min(passes, 39)];
    }
    fprintf(stderr, "long codes per 1000 probes: %.2f\n", 1000.0 * g_long / std::max(1l, g_steps));
    printf("%-60s chunks %ld passes avg %.2f max %d | cost(steps) avg %.0f max %ld | decodes/sub %.2f relink %ld\n",
           argv[1], st.chunks, (double)st.passes_sum / st.chunks, st.passes_max, (double)st.cost_sum / st.chunks,
           st.cost_max, (double)st.decodes / N, st.relink);
    if (getenv("HIST")) {
        for (int i = 0; i < 40; ++i)
            if (st.pass_hist[i]) printf("  passes %2d: %d\n", i, st.pass_hist[i]);
    }
    return 0;
}
