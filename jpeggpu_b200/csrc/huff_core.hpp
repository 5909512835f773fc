// huff_core.hpp — the bit-serial Huffman inner loop, shared by every kernel that walks the
// entropy stream (sync, link, write) and compiled for the host as well so that the LUT builder
// and the loop can be unit-tested on a machine without a GPU.
//
// Replaces src/decode_huffman.cu:167-394 (get_category / decode_next_symbol / decode_subsequence)
// and the two bit readers of src/decode_huffman_reader.hpp.  Differences by design:
//  * a symbol belongs to the subsequence in which it STARTS; the exit state records how many
//    bits it ran past the boundary (the reference stops before such a symbol and keeps a cache);
//  * one LUT probe yields code length + magnitude size + run, so the sync passes never extract
//    the magnitude bits at all;
//  * the bit reader is a two-word register window over big-endian words (destuff stores the
//    scan that way); the bit position is absolute and the funnel shift wraps it modulo 32, the
//    window slides by predicated moves and the following word is fetched unconditionally every
//    iteration from shared memory, so the symbol loop has no data-dependent branch besides the
//    rare long-code path and no load latency on its dependent chain except the LUT probe;
//  * the table pair of the current data unit and of the next one come from one 8-byte
//    shared-memory record per MCU slot, fetched in the shadow of the LUT probe.
//
// Memory access goes through a policy type `Mem` (shared memory on the device, plain arrays on
// the host):
//   lut(i)        -> uint16_t LUT pool entry
//   pair(c)       -> SlotPair of MCU slot c
//   slot_tbls(c)  -> dc table slot | ac table slot << 8   (slow path only)
//   slow(slot)    -> HuffSlow
//   natural(i)    -> zig-zag to raster index
//   word(i)       -> big-endian word i of the scan window the policy was built for
#ifndef JB2_HUFF_CORE_HPP_
#define JB2_HUFF_CORE_HPP_

#include "common.hpp"

namespace jb2 {

// Per MCU slot c: the tables of its data unit and what follows it.
struct SlotPair {
    uint32_t x; // ac table offset of slot c | dc table offset of slot next(c) << 16
    uint32_t y; // next(c) | dc table offset of slot c << 16
};

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
    // T.81 F.2.2.1 (src/decode_huffman.cu:196-200); size >= 1
    return bits < (1u << (size - 1)) ? (int)bits - (int)(1u << size) + 1 : (int)bits;
}

// (hi:lo) << (shift mod 32), upper word
JB2_HD uint32_t funnel_left_wrap(uint32_t hi, uint32_t lo, uint32_t shift)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, shift);
#else
    return (uint32_t)(((((uint64_t)hi << 32) | lo) << (shift & 31u)) >> 32);
#endif
}

// Decode the symbols that start in [p, end).  State: p bit position inside the policy's word
// window, q linear slot counter (q & 63 = zig-zag index), c MCU slot of the current data unit.
// WRITE: store the non-zero AC coefficients at out[(q >> 6) * 64 + natural[q & 63]] and every DC
// difference (zero or not) at dc_out[q >> 6], while q < q_limit.
template <bool WRITE, class Mem>
JB2_HD void decode_span(const Mem& M, uint32_t& p, uint32_t end, uint32_t& q, uint32_t& c,
                        int16_t* out, int16_t* dc_out, uint32_t q_limit)
{
    if (p >= end) return;
    uint32_t pp = p, qq = q, cc = c;
    uint32_t wnext = (pp >> 5) + 2u;
    uint32_t w0 = M.word(wnext - 2u), w1 = M.word(wnext - 1u);
    SlotPair pr = M.pair(cc);
    uint32_t tbl = (qq & 63u) == 0 ? (pr.y >> 16) : (pr.x & 0xffffu);

    while (pp < end) {
        if (WRITE && qq >= q_limit) break;
        const uint32_t nx = M.word(wnext); // not needed before the window slides
        pr = M.pair(cc);
        const uint32_t peek = funnel_left_wrap(w0, w1, pp);
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
            const bool is_dc = (qq & 63u) == 0;
            if ((size != 0 || is_dc) && pos < block_end) {
                int v = 0;
                if (size != 0) {
                    const uint32_t bits = (peek << (skip - size)) >> (32 - size);
                    v = extend_value(bits, size);
                }
                if (is_dc) dc_out[pos >> 6] = (int16_t)v;
                else out[(size_t)(pos & ~63u) + M.natural(pos & 63u)] = (int16_t)v;
            }
        }
        const uint32_t pn = pp + skip;
        const bool slide = ((pn ^ pp) & 32u) != 0;
        w0 = slide ? w1 : w0;
        w1 = slide ? nx : w1;
        wnext += slide ? 1u : 0u;
        pp = pn;
        // slot bookkeeping: a data unit ends on EOB / slot 63 (or overshoot on a garbage stream)
        const uint32_t qn = qq + adv;
        const bool done = qn >= block_end;
        qq = done ? block_end : qn;
        cc = done ? (pr.y & 0xffu) : cc;
        tbl = done ? (pr.x >> 16) : (pr.x & 0xffffu);
    }
    p = pp;
    q = qq;
    c = cc;
}

// Fills the per-slot records from the scan geometry (host and device).
JB2_HD void make_slot_pairs(const ScanGeom& g, SlotPair* pairs, uint32_t* tbls)
{
    for (int s = 0; s < kMaxDuPerMcu; ++s) {
        const int cur = s < g.dus_per_mcu ? s : 0;
        const int nxt = cur + 1 == g.dus_per_mcu ? 0 : cur + 1;
        pairs[s].x = (uint32_t)g.table_off[g.slot_ac[cur]] |
                     ((uint32_t)g.table_off[g.slot_dc[nxt]] << 16);
        pairs[s].y = (uint32_t)nxt | ((uint32_t)g.table_off[g.slot_dc[cur]] << 16);
        tbls[s] = (uint32_t)g.slot_dc[cur] | ((uint32_t)g.slot_ac[cur] << 8);
    }
}

// Plain-array policy (host emulation; also documents what the device policy must provide).
struct HostMem {
    const uint16_t* lut_;
    const SlotPair* pairs_;
    const uint32_t* tbls_; // per MCU slot: dc table slot | ac table slot << 8
    const HuffSlow* slow_;
    const uint8_t* natural_;
    const uint32_t* words_; // word 0 = first word of the restart segment
    JB2_HD uint16_t lut(uint32_t i) const { return lut_[i]; }
    JB2_HD SlotPair pair(uint32_t c) const { return pairs_[c]; }
    JB2_HD uint32_t slot_tbls(uint32_t c) const { return tbls_[c]; }
    JB2_HD const HuffSlow& slow(uint32_t s) const { return slow_[s]; }
    JB2_HD uint32_t natural(uint32_t i) const { return natural_[i]; }
    JB2_HD uint32_t word(uint32_t i) const { return words_[i]; }
};

} // namespace jb2
#endif
