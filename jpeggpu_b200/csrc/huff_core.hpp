// huff_core.hpp — the bit-serial Huffman inner loop, shared by every kernel that walks the
// entropy stream (speculate, repair, write) and compiled for the host as well so that the
// LUT builder and the loop can be unit-tested on a machine without a GPU.
//
// Replaces src/decode_huffman.cu:167-394 (get_category / decode_next_symbol / decode_subsequence)
// and the two bit readers of src/decode_huffman_reader.hpp.  Differences by design:
//  * a symbol belongs to the subsequence in which it STARTS; the exit state records how many
//    bits it ran past the boundary (the reference stops before such a symbol and keeps a cache);
//  * one LUT probe yields code length + magnitude size + run, so the sync passes never extract
//    the magnitude bits at all; the table for the next symbol is chosen when the slot
//    counter is updated, off the bit-position dependent chain;
//  * the bit buffer is a four-word register window read with one funnel shift per symbol and
//    refilled two words ahead by a predicated load (destuff stores the scan as big-endian
//    words), not a per-thread shared-memory cache: no branch and no exposed load latency
//    inside the symbol loop.
//
// Memory access goes through a policy type `Mem` (shared memory on the device, plain arrays on the
// host):  lut(i) -> uint16_t, slot_offs(c) -> dc_off | ac_off << 16, slot_tbls(c) -> dc table
// slot | ac table slot << 8, word(words, i) -> one big-endian word of the scan (word_if: predicated form), slow(slot) -> HuffSlow,
// natural(i) -> zig-zag to raster index.
#ifndef JB2_HUFF_CORE_HPP_
#define JB2_HUFF_CORE_HPP_

#include "common.hpp"

namespace jb2 {

// canonical search for codes the two LUT levels do not cover (pathological tables only)
JB2_HD uint16_t slow_entry(const HuffSlow& t, uint32_t peek16)
{
    int l = kLutBits + 1;
    for (; l < 16; ++l)
        if (peek16 < t.limit[l]) break;
    const uint32_t code = peek16 >> (16 - l);
    const uint8_t sym = t.huffval[(uint8_t)(t.valoff[l] + (int)code)];
    return make_entry(sym, l, t.is_dc != 0);
}

// second-level lookup for codes longer than kLutBits
template <class Mem>
JB2_HD uint16_t lookup_long(const Mem& M, uint32_t tbl_off, uint32_t tbl_slot, uint32_t peek16,
                            uint16_t e)
{
    if (e == kEntryLong) {
        const uint32_t base16 = M.lut(tbl_off - 8);
        const uint32_t n2 = M.lut(tbl_off - 7);
        const uint32_t i2 = peek16 - base16;
        e = i2 < n2 ? M.lut(tbl_off + kLut1Size + i2) : kEntrySlow;
    }
    if (e == kEntrySlow) e = slow_entry(M.slow(tbl_slot), peek16);
    return e;
}

JB2_HD int extend_value(uint32_t bits, uint32_t size)
{
    // T.81 F.2.2.1 (src/decode_huffman.cu:196-200)
    return bits < (1u << (size - 1)) ? (int)bits - (int)(1u << size) + 1 : (int)bits;
}

JB2_HD uint32_t funnel_left(uint32_t hi, uint32_t lo, uint32_t shift) // shift in [0, 31]
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, shift);
#else
    return (uint32_t)(((((uint64_t)hi << 32) | lo) << shift) >> 32);
#endif
}

// Per-scan table selection held in registers: the DU slots of an MCU are grouped by component
// (at most four groups), so the table pair of slot c is a three-compare select chain instead of
// a memory lookup (the reference does the same with an if-chain, src/decode_huffman.cu:323-338).
struct SlotTables {
    uint32_t t1, t2, t3;   // first slot of component group 1, 2, 3 (== K when absent)
    uint32_t o0, o1, o2, o3; // dc_off | ac_off << 16 per group
    uint32_t K;
    JB2_HD uint32_t offs(uint32_t c) const { return c < t1 ? o0 : (c < t2 ? o1 : (c < t3 ? o2 : o3)); }
};

template <class Mem>
JB2_HD SlotTables make_slot_tables(const Mem& M, uint32_t K)
{
    SlotTables S;
    S.K = K;
    S.o0 = S.o1 = S.o2 = S.o3 = M.slot_offs(0);
    S.t1 = S.t2 = S.t3 = K;
    uint32_t groups = 0, prev = S.o0;
    for (uint32_t c = 1; c < K; ++c) {
        const uint32_t v = M.slot_offs(c);
        if (v != prev && groups < 3) {
            ++groups;
            if (groups == 1) S.t1 = c, S.o1 = S.o2 = S.o3 = v;
            else if (groups == 2) S.t2 = c, S.o2 = S.o3 = v;
            else S.t3 = c, S.o3 = v;
            prev = v;
        }
    }
    return S;
}

// Decode symbols that start in [p, end).  State: p bit position (segment relative), q linear
// slot counter (q & 63 = zig-zag index), c DU index inside the MCU (K DUs per MCU).
// WRITE: store non-zero coefficients at out[(q >> 6) * 64 + natural[q & 63]] while q < q_limit.
//
// Bit reader: a sliding window of four big-endian words w0..w3; the next 32 bits of the stream
// are funnel(w0:w1 << o) with o in [0, 32).  Consuming bits only adds to o; when o crosses 32
// the window slides by one word and the word two ahead is fetched (predicated, no branch), so a
// load has at least 64 bits of decoding to complete before its value is needed.
// The loop body is branch-free apart from the rare long-code path: lanes of a warp sit at
// unrelated stream positions, so any data-dependent branch would be taken by some lane in
// nearly every iteration and serialise the whole warp.
template <bool WRITE, class Mem>
JB2_HD void decode_span(const Mem& M, const SlotTables& S, const uint32_t* words, uint32_t& p,
                        uint32_t end, uint32_t& q, uint32_t& c, int16_t* out, uint32_t q_limit)
{
    if (p >= end) return;
    uint32_t pp = p, qq = q, cc = c;
    uint32_t wi = pp >> 5;
    uint32_t o = pp & 31u;
    uint32_t w0 = M.word(words, wi), w1 = M.word(words, wi + 1), w2 = M.word(words, wi + 2),
             w3 = M.word(words, wi + 3);
    wi += 4;

    uint32_t offs = S.offs(cc);
    uint32_t tbl = (qq & 63u) == 0 ? (offs & 0xffffu) : (offs >> 16);

    while (pp < end) {
        if (WRITE && qq >= q_limit) break;
        const uint32_t peek = funnel_left(w0, w1, o);
        uint32_t e = M.lut(tbl + (peek >> (32 - kLutBits)));
        if (e >= kEntrySlow) {
            const uint32_t tbls = M.slot_tbls(cc);
            e = lookup_long(M, tbl, (qq & 63u) == 0 ? (tbls & 0xffu) : (tbls >> 8), peek >> 16,
                            (uint16_t)e);
        }
        const uint32_t skip = e & 31u;
        const uint32_t adv = e >> 9;
        const uint32_t block_end = (qq | 63u) + 1u;
        if (WRITE) {
            const uint32_t size = (e >> 5) & 15u;
            const uint32_t pos = qq + adv - 1u;
            if (size != 0 && pos < block_end) {
                const uint32_t bits = (peek << (skip - size)) >> (32 - size);
                out[(size_t)(pos & ~63u) + M.natural(pos & 63u)] = (int16_t)extend_value(bits, size);
            }
        }
        o += skip;
        pp += skip;
        const bool slide = o >= 32u;
        w0 = slide ? w1 : w0;
        w1 = slide ? w2 : w1;
        w2 = slide ? w3 : w2;
        M.word_if(words, wi, slide, w3); // predicated load straight into w3: not waited for here
        o = slide ? o - 32u : o;
        wi += slide ? 1u : 0u;
        // slot bookkeeping: a data unit ends on EOB / slot 63 (or overshoot on a garbage stream)
        const uint32_t qn = qq + adv;
        const bool done = qn >= block_end;
        qq = done ? block_end : qn;
        const uint32_t c1 = cc + 1u == S.K ? 0u : cc + 1u;
        cc = done ? c1 : cc;
        offs = S.offs(cc);
        tbl = done ? (offs & 0xffffu) : (offs >> 16);
    }
    p = pp;
    q = qq;
    c = cc;
}

// Plain-array policy (host emulation; also documents what the device policy must provide).
struct HostMem {
    const uint16_t* lut_;
    const uint32_t* offs_; // per DU slot: dc_off | ac_off << 16
    const uint32_t* tbls_; // per DU slot: dc table slot | ac table slot << 8
    const HuffSlow* slow_;
    const uint8_t* natural_;
    JB2_HD uint16_t lut(uint32_t i) const { return lut_[i]; }
    JB2_HD uint32_t slot_offs(uint32_t c) const { return offs_[c]; }
    JB2_HD uint32_t slot_tbls(uint32_t c) const { return tbls_[c]; }
    JB2_HD const HuffSlow& slow(uint32_t s) const { return slow_[s]; }
    JB2_HD uint32_t natural(uint32_t i) const { return natural_[i]; }
    JB2_HD uint32_t word(const uint32_t* words, uint32_t i) const { return words[i]; }
    JB2_HD void word_if(const uint32_t* words, uint32_t i, bool pred, uint32_t& dst) const
    {
        if (pred) dst = words[i];
    }
};

} // namespace jb2
#endif
