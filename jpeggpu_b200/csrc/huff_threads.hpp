// huff_threads.hpp — the per-thread bodies of the Huffman sync / link / write kernels as
// host+device functions, so tests/host_emul can run the exact same logic sequentially on a CPU.
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
JB2_HD uint16_t entry_g(unsigned long long e) { return (uint16_t)e; }
JB2_HD uint16_t entry_x(unsigned long long e) { return (uint16_t)(e >> 16); }
JB2_HD uint32_t entry_n(unsigned long long e) { return (uint32_t)(e >> 32); }

JB2_HD uint32_t subs_in_segment(const SegDesc& seg)
{
    return (seg.nbytes + (uint32_t)kSubBytes - 1u) / (uint32_t)kSubBytes;
}

// Decode subsequence i of a segment (policy M addresses words relative to the segment start)
// from packed entry state `gs` (15 bits); returns {gs, exit state, slots advanced} and, in *mid,
// the state at the middle of the subsequence | slots advanced before it << 16 (the write pass
// splits every subsequence between two threads there).
template <class Mem>
JB2_HD unsigned long long decode_sub(const Mem& M, uint32_t seg_bits, uint32_t i, uint16_t gs,
                                     uint32_t* mid)
{
    const uint32_t start = i * (uint32_t)kSubBits;
    const uint32_t half = start + (uint32_t)kSubBits / 2u;
    const uint32_t sub_end = start + (uint32_t)kSubBits;
    uint32_t p = start + (gs & 31u);
    uint32_t c = (gs >> 5) & 15u;
    const uint32_t z = (gs >> 9) & 63u;
    uint32_t q = z;
    decode_span<false>(M, p, umin(half, seg_bits), q, c, 0u);
    *mid = pack_state(p > half ? p - half : 0u, c, q & 63u) | ((q - z) << 16);
    decode_span<false>(M, p, umin(sub_end, seg_bits), q, c, 0u);
    const uint32_t o = p > sub_end ? p - sub_end : 0u;
    return pack_entry(gs, pack_state(o, c, q & 63u), q - z);
}

// huff_link: does the record of subsequence j start where its predecessor stopped?
JB2_HD bool boundary_ok(unsigned long long prev, unsigned long long cur)
{
    return (entry_g(cur) & kStateHead) != 0 || (entry_g(cur) & kStateMask) == entry_x(prev);
}

// huff_link fallback: re-decode forward from subsequence j (entry = exit of j-1) until the
// stored records agree again.  M addresses words relative to the segment start; `lo` / `hi` are
// the two words of every record (stride 2).  Returns the number of records rewritten.
template <class Mem, typename WordPtr>
JB2_HD uint32_t relink_from(const Mem& M, const SegDesc& seg, WordPtr rec_words, uint32_t j)
{
    const uint32_t seg_bits = seg.nbytes * 8u;
    const uint32_t seg_end = seg.sub_off + subs_in_segment(seg);
    uint16_t state = entry_x(rec_words[2 * (size_t)(j - 1)]);
    uint32_t rewritten = 0;
    for (uint32_t m = j; m < seg_end; ++m) {
        if (m != j && (entry_g(rec_words[2 * (size_t)m]) & kStateMask) == state) break;
        uint32_t mid;
        const unsigned long long e = decode_sub(M, seg_bits, m - seg.sub_off, state, &mid);
        rec_words[2 * (size_t)m + 1] = mid;
        rec_words[2 * (size_t)m] = e;
        state = entry_x(e);
        ++rewritten;
    }
    return rewritten;
}

// Placement of a restart segment in the coefficient buffer (T.81 E.1.4: every segment but the
// last holds Ri MCUs).  Returns false for a surplus segment (nothing to place).
JB2_HD bool segment_extent(const ScanGeom& g, uint32_t k, uint32_t& first_du, uint32_t& q_limit)
{
    const unsigned long long mcu0 = (unsigned long long)k * (unsigned)g.restart_interval;
    if (mcu0 >= (unsigned long long)g.total_mcus) return false;
    const uint32_t mcus = umin((uint32_t)g.restart_interval, (uint32_t)g.total_mcus - (uint32_t)mcu0);
    first_du = (uint32_t)mcu0 * (uint32_t)g.dus_per_mcu;
    q_limit = mcus * (uint32_t)g.dus_per_mcu * 64u;
    return true;
}

// huff_write body: decode one half (h = 0, 1) of subsequence i from its verified record and hand
// the coefficients to the policy.  q0 = slot (segment relative) at which the subsequence starts.
template <class Mem>
JB2_HD void write_half(const Mem& M, const SegDesc& seg, uint32_t i, uint32_t h,
                       unsigned long long rec, uint32_t mid, uint32_t q0, uint32_t q_limit)
{
    const uint32_t start = i * (uint32_t)kSubBits;
    const uint32_t half = start + (uint32_t)kSubBits / 2u;
    const uint32_t seg_bits = seg.nbytes * 8u;
    const uint32_t gs = h ? (mid & 0x7fffu) : (entry_g(rec) & kStateMask);
    uint32_t q = h ? q0 + (mid >> 16) : q0;
    uint32_t c = (gs >> 5) & 15u;
    uint32_t p = (h ? half : start) + (gs & 31u);
    const uint32_t end = umin(h ? start + (uint32_t)kSubBits : half, seg_bits);
    decode_span<true>(M, p, end, q, c, q_limit);
}

} // namespace jb2
#endif
