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
// from packed entry state `gs` (15 bits); returns {gs, exit state, slots advanced}.
template <class Mem>
JB2_HD unsigned long long decode_sub(const Mem& M, uint32_t seg_bits, uint32_t i, uint16_t gs)
{
    const uint32_t start = i * (uint32_t)kSubBits;
    const uint32_t sub_end = start + (uint32_t)kSubBits;
    const uint32_t end = umin(sub_end, seg_bits);
    uint32_t p = start + (gs & 31u);
    uint32_t c = (gs >> 5) & 15u;
    const uint32_t z = (gs >> 9) & 63u;
    uint32_t q = z;
    decode_span<false>(M, p, end, q, c, nullptr, nullptr, 0u);
    const uint32_t o = p > sub_end ? p - sub_end : 0u;
    return pack_entry(gs, pack_state(o, c, q & 63u), q - z);
}

// huff_link: does the record of subsequence j start where its predecessor stopped?
JB2_HD bool boundary_ok(unsigned long long prev, unsigned long long cur)
{
    return (entry_g(cur) & kStateHead) != 0 || (entry_g(cur) & kStateMask) == entry_x(prev);
}

// huff_link fallback: re-decode forward from subsequence j (entry = exit of j-1) until the
// stored records agree again.  M addresses words relative to the segment start.  Returns the
// number of records rewritten.
template <class Mem, typename EntryPtr>
JB2_HD uint32_t relink_from(const Mem& M, const SegDesc& seg, EntryPtr entries, uint32_t j)
{
    const uint32_t seg_bits = seg.nbytes * 8u;
    const uint32_t seg_end = seg.sub_off + subs_in_segment(seg);
    uint16_t state = entry_x(entries[j - 1]);
    uint32_t rewritten = 0;
    for (uint32_t m = j; m < seg_end; ++m) {
        if (m != j && (entry_g(entries[m]) & kStateMask) == state) break;
        const unsigned long long e = decode_sub(M, seg_bits, m - seg.sub_off, state);
        entries[m] = e;
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

// huff_write body: decode subsequence i of segment k from its verified record and place the
// coefficients.  q0 = slot (segment relative) at which the subsequence starts.
template <class Mem>
JB2_HD void write_sub(const Mem& M, const ScanGeom& g, const SegDesc& seg, uint32_t k, uint32_t i,
                      uint16_t gs, uint32_t q0, int16_t* coef, int16_t* dc)
{
    uint32_t first_du, q_limit;
    if (!segment_extent(g, k, first_du, q_limit)) return;
    uint32_t q = q0;
    uint32_t c = (gs >> 5) & 15u;
    uint32_t p = i * (uint32_t)kSubBits + (gs & 31u);
    const uint32_t end = umin((i + 1u) * (uint32_t)kSubBits, seg.nbytes * 8u);
    decode_span<true>(M, p, end, q, c, coef + (size_t)first_du * 64, dc + first_du, q_limit);
}

} // namespace jb2
#endif
