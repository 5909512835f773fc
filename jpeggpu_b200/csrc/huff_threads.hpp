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
// from packed entry state `gs` (15 bits); returns {gs, exit state, slots advanced} and, in *mids,
// the state at every inner part boundary | slots advanced before it << 16 (the write pass
// splits every subsequence between kWriteParts threads there).
template <class Mem>
JB2_HD unsigned long long decode_sub(const Mem& M, uint32_t seg_bits, uint32_t i, uint16_t gs,
                                     Mids* mids)
{
    const uint32_t start = i * (uint32_t)kSubBits;
    const uint32_t sub_end = start + (uint32_t)kSubBits;
    uint32_t p = start + (gs & 31u);
    uint32_t c = (gs >> 5) & 15u;
    const uint32_t z = (gs >> 9) & 63u;
    uint32_t q = z;
#pragma unroll
    for (int k = 1; k < kWriteParts; ++k) {
        const uint32_t edge = start + (uint32_t)(k * (kSubBits / kWriteParts));
        decode_span<false>(M, p, umin(edge, seg_bits), q, c, 0u);
        mids->v[k - 1] = pack_state(p > edge ? p - edge : 0u, c, q & 63u) | ((q - z) << 16);
    }
    decode_span<false>(M, p, umin(sub_end, seg_bits), q, c, 0u);
    const uint32_t o = p > sub_end ? p - sub_end : 0u;
    return pack_entry(gs, pack_state(o, c, q & 63u), q - z);
}

// The 64-bit words 1.. of a record
JB2_HD unsigned long long mids_word(const Mids& m, int w)
{
    const int k = 2 * w;
    unsigned long long v = m.v[k];
    if (k + 1 < kWriteParts - 1) v |= (unsigned long long)m.v[k + 1] << 32;
    return v;
}
JB2_HD uint32_t mid_of(const unsigned long long* rec_words, int k) // rec_words: words 1.. of a record
{
    return (uint32_t)(rec_words[k >> 1] >> (32 * (k & 1)));
}

// verification (last CTA of huff_sync): does the record of subsequence j start where its
// predecessor stopped?
JB2_HD bool boundary_ok(unsigned long long prev, unsigned long long cur)
{
    return (entry_g(cur) & kStateHead) != 0 || (entry_g(cur) & kStateMask) == entry_x(prev);
}

// Stores the record of subsequence m.  Two-word records go out as ONE 16-byte store on the device:
// verify_and_repair lets several threads repair overlapping ranges, and a record assembled from
// two writers (entry word of one, mid states of another) would pass the boundary check.
template <typename WordPtr>
JB2_HD void store_record(WordPtr rec_words, size_t m, unsigned long long e, const Mids& mids)
{
#if defined(__CUDA_ARCH__)
    if (kRecWords == 2) {
        asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(rec_words + kRecWords * m), "l"(e),
                     "l"(mids_word(mids, 0))
                     : "memory");
        return;
    }
#endif
    for (int w = 0; 2 * w < kWriteParts - 1; ++w) rec_words[kRecWords * m + 1 + w] = mids_word(mids, w);
    rec_words[kRecWords * m] = e;
}

// repair: re-decode forward from subsequence j (entry = exit of j-1) until the
// stored records agree again.  M addresses words relative to the segment start; rec_words holds
// kRecWords 64-bit words per record.  Returns the number of records rewritten.
template <class Mem, typename WordPtr>
JB2_HD uint32_t relink_from(const Mem& M, const SegDesc& seg, WordPtr rec_words, uint32_t j)
{
    const uint32_t seg_bits = seg.nbytes * 8u;
    const uint32_t seg_end = seg.sub_off + subs_in_segment(seg);
    uint16_t state = entry_x(rec_words[kRecWords * (size_t)(j - 1)]);
    uint32_t rewritten = 0;
    for (uint32_t m = j; m < seg_end; ++m) {
        if (m != j && (entry_g(rec_words[kRecWords * (size_t)m]) & kStateMask) == state) break;
        Mids mids;
        const unsigned long long e = decode_sub(M, seg_bits, m - seg.sub_off, state, &mids);
        store_record(rec_words, (size_t)m, e, mids);
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

// huff_write body: decode part h (0 .. kWriteParts-1) of subsequence i from its verified record
// and hand the coefficients to the policy.  q0 = slot (segment relative) at which the
// subsequence starts; mid = the record's state at the part's start (ignored for h == 0).
template <class Mem>
JB2_HD void write_part(const Mem& M, const SegDesc& seg, uint32_t i, uint32_t h,
                       unsigned long long rec, uint32_t mid, uint32_t q0, uint32_t q_limit)
{
    const uint32_t part_bits = (uint32_t)(kSubBits / kWriteParts);
    const uint32_t begin = i * (uint32_t)kSubBits + h * part_bits;
    const uint32_t seg_bits = seg.nbytes * 8u;
    const uint32_t gs = h ? (mid & 0x7fffu) : (entry_g(rec) & kStateMask);
    uint32_t q = h ? q0 + (mid >> 16) : q0;
    uint32_t c = (gs >> 5) & 15u;
    uint32_t p = begin + (gs & 31u);
    const uint32_t end = umin(begin + part_bits, seg_bits);
    decode_span<true>(M, p, end, q, c, q_limit);
}

} // namespace jb2
#endif
