// huff_core.hpp — the bit-serial Huffman inner loop, shared by every kernel that walks the
// entropy stream (speculate, repair, write) and compiled for the host as well so that the
// LUT builder and the loop can be unit-tested on a machine without a GPU.
//
// Replaces src/decode_huffman.cu:167-394 (get_category / decode_next_symbol / decode_subsequence)
// and the two bit readers of src/decode_huffman_reader.hpp.  Differences by design:
//  * a symbol belongs to the subsequence in which it STARTS; the exit state records how many
//    bits it ran past the boundary (the reference stops before such a symbol and keeps a cache);
//  * one LUT probe yields code length + magnitude size + run, so the sync passes never extract
//    the magnitude bits at all; the DC and the AC table are probed in parallel and selected
//    afterwards, which keeps the table choice off the dependent chain;
//  * the bit buffer is a 64-bit register fed from a 16-byte register chunk with the following
//    chunk already in flight (destuff stores the scan as big-endian words), not a per-thread
//    shared-memory cache: no global-load latency is exposed inside the symbol loop.
//
// Memory access goes through a policy type `Mem` (shared memory on the device, plain arrays on the
// host):  lut(i) -> uint16_t, slot_offs(c) -> dc_off | ac_off << 16, slot_tbls(c) -> dc table
// slot | ac table slot << 8, chunk(words, i) -> 4 consecutive words, slow(slot) -> HuffSlow,
// natural(i) -> zig-zag to raster index.
#ifndef JB2_HUFF_CORE_HPP_
#define JB2_HUFF_CORE_HPP_

#include "common.hpp"

namespace jb2 {

struct Chunk4 {
    uint32_t x, y, z, w;
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
    // T.81 F.2.2.1 (src/decode_huffman.cu:196-200)
    return bits < (1u << (size - 1)) ? (int)bits - (int)(1u << size) + 1 : (int)bits;
}

// Decode symbols that start in [p, end).  State: p bit position (segment relative), q linear
// slot counter (q & 63 = zig-zag index), c DU index inside the MCU (K DUs per MCU).
// WRITE: store non-zero coefficients at out[(q >> 6) * 64 + natural[q & 63]] while q < q_limit.
template <bool WRITE, class Mem>
JB2_HD void decode_span(const Mem& M, uint32_t K, const uint32_t* words, uint32_t& p,
                        uint32_t end, uint32_t& q, uint32_t& c, int16_t* out, uint32_t q_limit)
{
    if (p >= end) return;
    uint32_t pp = p, qq = q, cc = c;

    // bit reader: `cur` is the chunk being consumed, `nxt` the prefetched one
    uint32_t ci = pp >> 7;
    uint32_t k = (pp >> 5) & 3u;
    Chunk4 cur = M.chunk(words, ci), nxt = M.chunk(words, ci + 1);
    auto take = [&]() -> uint32_t {
        const uint32_t w = k == 0 ? cur.x : (k == 1 ? cur.y : (k == 2 ? cur.z : cur.w));
        if (++k == 4) {
            cur = nxt;
            ++ci;
            nxt = M.chunk(words, ci + 1);
            k = 0;
        }
        return w;
    };
    const uint32_t w0 = take(), w1 = take();
    uint64_t buf = (((uint64_t)w0 << 32) | (uint64_t)w1) << (pp & 31u);
    int avail = 64 - (int)(pp & 31u);

    uint32_t offs = M.slot_offs(cc);
    uint32_t cn = cc + 1 == K ? 0 : cc + 1;
    uint32_t offs_next = M.slot_offs(cn);

    while (pp < end) {
        if (WRITE && qq >= q_limit) break;
        if (avail <= 32) {
            buf |= (uint64_t)take() << (32 - avail);
            avail += 32;
        }
        const uint32_t hi = (uint32_t)(buf >> 32);
        const uint32_t i1 = hi >> (32 - kLutBits);
        const uint16_t e_dc = M.lut((offs & 0xffffu) + i1);
        const uint16_t e_ac = M.lut((offs >> 16) + i1);
        const bool is_dc = (qq & 63u) == 0;
        uint16_t e = is_dc ? e_dc : e_ac;
        if (e >= kEntrySlow) {
            const uint32_t tbls = M.slot_tbls(cc);
            e = lookup_long(M, is_dc ? (offs & 0xffffu) : (offs >> 16),
                            is_dc ? (tbls & 0xffu) : (tbls >> 8), hi >> 16, e);
        }
        const uint32_t skip = e & 31u;
        const uint32_t size = (e >> 5) & 15u;
        const uint32_t adv = e >> 9;
        const uint32_t block_end = (qq | 63u) + 1u;
        if (WRITE) {
            const uint32_t pos = qq + adv - 1u;
            if (size != 0 && pos < block_end) {
                const uint32_t bits = (uint32_t)((buf << (skip - size)) >> (64 - size));
                out[(size_t)(pos & ~63u) + M.natural(pos & 63u)] = (int16_t)extend_value(bits, size);
            }
        }
        buf <<= skip;
        avail -= (int)skip;
        pp += skip;
        uint32_t qn = qq + adv;
        if (qn >= block_end) { // data unit complete (or overshoot on a garbage stream)
            qn = block_end;
            cc = cn;
            offs = offs_next;
            cn = cc + 1 == K ? 0 : cc + 1;
            offs_next = M.slot_offs(cn);
        }
        qq = qn;
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
    JB2_HD Chunk4 chunk(const uint32_t* words, uint32_t i) const
    {
        const uint32_t* w = words + (size_t)i * 4;
        return Chunk4{w[0], w[1], w[2], w[3]};
    }
};

} // namespace jb2
#endif
