// huff_threads.hpp — the per-thread bodies of the Huffman sync / write kernels as host+device
// functions, so tests/host_emul can run the exact same logic sequentially on a CPU.
#ifndef JB2_HUFF_THREADS_HPP_
#define JB2_HUFF_THREADS_HPP_

#include "huff_core.hpp"

namespace jb2 {

JB2_HD uint32_t umin(uint32_t a, uint32_t b) { return a < b ? a : b; }

// last segment whose first subsequence is <= j (empty segments share their successor's offset)
JB2_HD uint32_t find_segment(const SegDesc* segs, int n, uint32_t j)
{
    int lo = 0, hi = n; // invariant: segs[lo].sub_off <= j, answer in [lo, hi)
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (segs[mid].sub_off <= j) lo = mid;
        else hi = mid;
    }
    return (uint32_t)lo;
}

JB2_HD unsigned long long pack_entry(uint16_t gs, uint16_t xs, uint32_t n)
{
    return (unsigned long long)gs | ((unsigned long long)xs << 16) | ((unsigned long long)n << 32);
}

// Decode subsequence i of a segment from packed entry state `gs`; returns {gs, exit, n}.
template <class Mem>
JB2_HD unsigned long long decode_sub(const Mem& M, const SlotTables& K, const uint32_t* words,
                                     uint32_t seg_bits, uint32_t i, uint16_t gs)
{
    const uint32_t start = i * (uint32_t)kSubBits;
    const uint32_t sub_end = start + (uint32_t)kSubBits;
    const uint32_t end = umin(sub_end, seg_bits);
    uint32_t p = start + (gs & 31u);
    uint32_t c = (gs >> 5) & 15u;
    const uint32_t z = gs >> 9;
    uint32_t q = z;
    decode_span<false>(M, K, words, p, end, q, c, nullptr, 0u);
    const uint32_t o = p > sub_end ? p - sub_end : 0u;
    return pack_entry(gs, pack_state(o, c, q & 63u), q - z);
}

// huff_speculate body: guess the entry state of subsequence j by a cold run-in, then decode it.
template <class Mem>
JB2_HD unsigned long long speculate_one(const Mem& M, const SlotTables& K, const SegDesc* segs, int num_segs,
                                        const uint32_t* destuffed, uint32_t j, uint32_t* seg_out)
{
    const uint32_t k = find_segment(segs, num_segs, j);
    const SegDesc seg = segs[k];
    *seg_out = k;
    const uint32_t i = j - seg.sub_off;
    const uint32_t* words = destuffed + (size_t)seg.sub_off * kSubWords;
    // run-in: cold start (or the true segment start) up to kRunInSubs subsequences earlier
    const uint32_t i0 = i > (uint32_t)kRunInSubs ? i - (uint32_t)kRunInSubs : 0u;
    uint32_t p = i0 * (uint32_t)kSubBits, q = 0, c = 0;
    const uint32_t boundary = i * (uint32_t)kSubBits;
    decode_span<false>(M, K, words, p, boundary, q, c, nullptr, 0u);
    const uint16_t gs = pack_state(p - boundary, c, q & 63u);
    return decode_sub(M, K, words, seg.nbytes * 8u, i, gs);
}

// huff_repair body. Returns the number of records rewritten.
template <class Mem, typename EntryPtr>
JB2_HD uint32_t repair_one(const Mem& M, const SlotTables& K, const SegDesc* segs, const uint32_t* seg_idx,
                           const uint32_t* destuffed, EntryPtr entries, uint32_t j)
{
    const SegDesc seg = segs[seg_idx[j]];
    if (j == seg.sub_off) return 0; // first record of a segment starts from the initial state
    uint16_t state = (uint16_t)(entries[j - 1] >> 16);
    if (state == (uint16_t)entries[j]) return 0;
    const uint32_t* words = destuffed + (size_t)seg.sub_off * kSubWords;
    const uint32_t seg_bits = seg.nbytes * 8u;
    const uint32_t seg_end = seg.sub_off + (seg.nbytes + kSubBytes - 1) / kSubBytes;
    uint32_t repaired = 0;
    for (uint32_t m = j; m < seg_end; ++m) {
        if (m != j && state == (uint16_t)entries[m]) break;
        const unsigned long long e = decode_sub(M, K, words, seg_bits, m - seg.sub_off, state);
        entries[m] = e;
        state = (uint16_t)(e >> 16);
        ++repaired;
    }
    return repaired;
}

// huff_fixup body: sequential verification of one segment.
template <class Mem>
JB2_HD void fixup_segment(const Mem& M, const SlotTables& K, const SegDesc seg, const uint32_t* destuffed,
                          unsigned long long* all_entries)
{
    const uint32_t* words = destuffed + (size_t)seg.sub_off * kSubWords;
    const uint32_t seg_bits = seg.nbytes * 8u;
    const uint32_t count = (seg.nbytes + kSubBytes - 1) / kSubBytes;
    unsigned long long* entries = all_entries + seg.sub_off;
    if (count == 0) return;
    uint16_t state = (uint16_t)(entries[0] >> 16);
    for (uint32_t i = 1; i < count; ++i) {
        unsigned long long e = entries[i];
        if ((uint16_t)e != state) {
            e = decode_sub(M, K, words, seg_bits, i, state);
            entries[i] = e;
        }
        state = (uint16_t)(e >> 16);
    }
}

// huff_write body.
template <class Mem>
JB2_HD void write_one(const Mem& M, const SlotTables& K, const ScanGeom& g, const SegDesc* segs,
                      const uint32_t* seg_idx, const uint32_t* destuffed,
                      const unsigned long long* entries, const uint32_t* qpos, int16_t* coef,
                      uint32_t j)
{
    const uint32_t k = seg_idx[j];
    const SegDesc seg = segs[k];
    // MCUs coded by this segment (T.81 E.1.4: every segment but the last holds Ri MCUs)
    const unsigned long long mcu0 = (unsigned long long)k * (unsigned)g.restart_interval;
    if (mcu0 >= (unsigned long long)g.total_mcus) return; // surplus segment: nothing to place
    const uint32_t mcus = umin((uint32_t)g.restart_interval, (uint32_t)g.total_mcus - (uint32_t)mcu0);
    const uint32_t q_limit = mcus * (uint32_t)g.dus_per_mcu * 64u;
    int16_t* out = coef + (size_t)mcu0 * g.dus_per_mcu * 64;
    const uint32_t i = j - seg.sub_off;
    const uint16_t gs = (uint16_t)entries[j];
    uint32_t q = qpos[j] - qpos[seg.sub_off];
    uint32_t c = (gs >> 5) & 15u;
    uint32_t p = i * (uint32_t)kSubBits + (gs & 31u);
    const uint32_t end = umin((i + 1u) * (uint32_t)kSubBits, seg.nbytes * 8u);
    const uint32_t* words = destuffed + (size_t)seg.sub_off * kSubWords;
    decode_span<true>(M, K, words, p, end, q, c, out, q_limit);
}

} // namespace jb2
#endif
