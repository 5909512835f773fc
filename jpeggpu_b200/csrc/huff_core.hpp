// huff_core.hpp — the bit-serial Huffman inner loop, shared by every kernel that walks the
// entropy stream (speculate, repair, write) and compiled for the host as well so that the
// LUT builder and the loop can be unit-tested on a machine without a GPU.
//
// Replaces src/decode_huffman.cu:167-394 (get_category / decode_next_symbol / decode_subsequence)
// and the two bit readers of src/decode_huffman_reader.hpp.  Differences by design:
//  * a symbol belongs to the subsequence in which it STARTS; the exit state records how many
//    bits it ran past the boundary (the reference stops before such a symbol and keeps a cache);
//  * one LUT probe yields code length + magnitude size + run, so the sync passes never extract
//    the magnitude bits at all;
//  * the bit buffer is a 64-bit register fed by 32-bit big-endian words (destuff stores the scan
//    already word-swapped), not a per-thread shared-memory cache.
#ifndef JB2_HUFF_CORE_HPP_
#define JB2_HUFF_CORE_HPP_

#include "common.hpp"

namespace jb2 {

struct DecodeTables {
    const uint16_t* lut;    // LUT pool (shared memory on the device)
    const HuffSlow* slow;   // per table slot, global memory
    const uint16_t* dc_off; // per DU slot: u16 offset of the DC table's primary LUT
    const uint16_t* ac_off; // per DU slot: same for AC
    const uint8_t* dc_slot; // per DU slot: table slot (for the slow path)
    const uint8_t* ac_slot;
    int dus_per_mcu;
};

struct BitCursor {
    const uint32_t* words; // segment base (big-endian words)
    uint64_t buf;          // next bits, MSB first
    int avail;             // valid bits in buf (> 32 whenever a symbol is decoded)
    uint32_t next;         // index of the next word to append
};

JB2_HD void cursor_init(BitCursor& bc, const uint32_t* words, uint32_t bitpos)
{
    const uint32_t wi = bitpos >> 5;
    const int b = (int)(bitpos & 31u);
    bc.words = words;
    bc.buf = (((uint64_t)words[wi] << 32) | (uint64_t)words[wi + 1]) << b;
    bc.avail = 64 - b;
    bc.next = wi + 2;
}

JB2_HD void cursor_refill(BitCursor& bc)
{
    if (bc.avail <= 32) {
        bc.buf |= (uint64_t)bc.words[bc.next] << (32 - bc.avail);
        bc.avail += 32;
        ++bc.next;
    }
}

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

JB2_HD uint16_t lookup(const DecodeTables& T, uint32_t tbl_off, uint32_t tbl_slot,
                       uint32_t peek16)
{
    uint16_t e = T.lut[tbl_off + (peek16 >> (16 - kLutBits))];
    if (e >= kEntrySlow) {
        if (e == kEntryLong) {
            const uint32_t base16 = T.lut[tbl_off - 8];
            const uint32_t n2 = T.lut[tbl_off - 7];
            const uint32_t i2 = peek16 - base16;
            e = i2 < n2 ? T.lut[tbl_off + kLut1Size + i2] : kEntrySlow;
        }
        if (e == kEntrySlow) e = slow_entry(T.slow[tbl_slot], peek16);
    }
    return e;
}

JB2_HD int extend_value(uint32_t bits, uint32_t size)
{
    // T.81 F.2.2.1 (src/decode_huffman.cu:196-200)
    return bits < (1u << (size - 1)) ? (int)bits - (int)(1u << size) + 1 : (int)bits;
}

// Decode symbols that start in [p, end).  State: p bit position (segment relative), q linear
// slot counter (q & 63 = zig-zag index), c DU index inside the MCU.
// WRITE: store non-zero coefficients at out[(q >> 6) * 64 + natural[q & 63]] while q < q_limit.
template <bool WRITE>
JB2_HD void decode_span(const DecodeTables& T, const uint32_t* words, uint32_t& p, uint32_t end,
                        uint32_t& q, uint32_t& c, int16_t* out, uint32_t q_limit,
                        const uint8_t* natural)
{
    if (p >= end) return;
    BitCursor bc;
    cursor_init(bc, words, p);
    uint32_t dc_off = T.dc_off[c], ac_off = T.ac_off[c];
    uint32_t pp = p, qq = q, cc = c;
    while (pp < end) {
        if (WRITE && qq >= q_limit) break;
        cursor_refill(bc);
        const uint32_t peek16 = (uint32_t)(bc.buf >> 48);
        const bool is_dc = (qq & 63u) == 0;
        const uint16_t e = lookup(T, is_dc ? dc_off : ac_off,
                                  is_dc ? T.dc_slot[cc] : T.ac_slot[cc], peek16);
        const uint32_t skip = e & 31u;
        const uint32_t size = (e >> 5) & 15u;
        const uint32_t adv = e >> 9;
        const uint32_t block_end = (qq | 63u) + 1u;
        if (WRITE) {
            const uint32_t pos = qq + adv - 1u;
            if (size != 0 && pos < block_end) {
                const uint32_t bits = (uint32_t)((bc.buf << (skip - size)) >> (64 - size));
                out[(size_t)(pos & ~63u) + natural[pos & 63u]] = (int16_t)extend_value(bits, size);
            }
        }
        bc.buf <<= skip;
        bc.avail -= (int)skip;
        pp += skip;
        uint32_t qn = qq + adv;
        if (qn >= block_end) { // data unit complete (or overshoot on a garbage stream)
            qn = block_end;
            cc = cc + 1 == (uint32_t)T.dus_per_mcu ? 0 : cc + 1;
            dc_off = T.dc_off[cc];
            ac_off = T.ac_off[cc];
        }
        qq = qn;
    }
    p = pp;
    q = qq;
    c = cc;
}

} // namespace jb2
#endif
