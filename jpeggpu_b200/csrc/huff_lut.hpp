// huff_lut.hpp — host-side expansion of a DHT specification into the two-level decode LUT
// (layout in common.hpp).  Replaces compute_huffman_table of src/reader.cpp:186-224, whose
// 8-bit {value, nbits} LUT + maxcode search becomes a 10-bit primary + exact secondary LUT that
// already folds the magnitude size and the run into one 32-bit entry.
#ifndef JB2_HUFF_LUT_HPP_
#define JB2_HUFF_LUT_HPP_

#include "common.hpp"
#include "parser.hpp"

#include <vector>

namespace jb2 {

// Appends one table to `pool`; returns the u16 offset of its primary LUT. Fills `slow`.
inline uint32_t build_lut(const HuffSpec& spec, std::vector<uint32_t>& pool, HuffSlow& slow)
{
    const bool is_dc = spec.is_dc;
    // canonical codes (T.81 Annex C)
    uint16_t code_of[256];
    uint8_t len_of[256];
    int nsym = 0;
    unsigned code = 0;
    for (int l = 1; l <= 16; ++l) {
        slow.limit[l] = 0;
        slow.valoff[l] = 0;
        if (spec.counts[l - 1]) {
            slow.valoff[l] = nsym - (int)code;
            for (int i = 0; i < spec.counts[l - 1]; ++i) {
                code_of[nsym] = (uint16_t)code++;
                len_of[nsym] = (uint8_t)l;
                ++nsym;
            }
            slow.limit[l] = code << (16 - l);
        }
        code <<= 1;
    }
    slow.limit[0] = 0;
    slow.valoff[0] = 0;
    // a length with no codes must never match in slow_entry: limit 0 does that (peek16 < 0 false)
    for (int i = 0; i < 256; ++i) slow.huffval[i] = i < nsym ? spec.vals[i] : 0;
    slow.is_dc = is_dc ? 1 : 0;
    slow.pad[0] = slow.pad[1] = slow.pad[2] = 0;

    // the long-code region is contiguous at the top of the 16-bit code space (canonical order)
    uint32_t base16 = 0x10000u;
    for (int i = 0; i < nsym; ++i)
        if (len_of[i] > kLutBits) {
            const uint32_t first = (uint32_t)code_of[i] << (16 - len_of[i]);
            const uint32_t aligned = first & ~((1u << (16 - kLutBits)) - 1u);
            if (aligned < base16) base16 = aligned;
        }
    uint32_t lut2_size = 0x10000u - base16;
    const bool truncated = lut2_size > (uint32_t)kLut2Max; // pathological: canonical search instead
    if (truncated) lut2_size = 0;
    const uint32_t lut2_alloc = (lut2_size + 3u) & ~3u; // every piece stays 16-byte aligned
    const uint32_t lut2_off = is_dc ? (uint32_t)kLut1Size : 2u * (uint32_t)kLut1Size;

    const size_t header_at = pool.size();
    pool.resize(header_at + kLutHeader + lut2_off + lut2_alloc);
    uint32_t* hdr = pool.data() + header_at;
    uint32_t* lut1 = hdr + kLutHeader;
    uint32_t* lut2 = lut1 + lut2_off;
    for (int i = 0; i < kLutHeader; ++i) hdr[i] = 0;
    hdr[0] = base16; // 0x10000 with lut2_size 0: never matches
    hdr[1] = lut2_size;
    hdr[2] = lut2_off;
    hdr[3] = is_dc ? 0u : 1u;
    // unused code space: behave like the reference's forced 16-bit termination
    // (src/decode_huffman.cu:177-193) with symbol 0
    const uint32_t invalid = make_entry(0, 16, is_dc);
    for (int i = 0; i < kLut1Size; ++i) lut1[i] = invalid;
    // secondary entries are reached through a continuation entry that has already consumed
    // kContBits bits of the code
    const uint32_t consumed = (uint32_t)kContBits << 27;
    for (uint32_t i = 0; i < lut2_alloc; ++i) lut2[i] = invalid - consumed;
    const uint32_t first_long_prefix = base16 >> (16 - kLutBits);
    // where a continuation sends the next probe: index i of that probe stands for the 16-bit
    // prefix kContBase16 + i, the entry of prefix base16 is lut2[0]
    const uint32_t lut2_word = (uint32_t)(header_at + kLutHeader) + lut2_off; // pool word offset of lut2[0]
    const uint32_t cont = truncated ? kEntrySlow
                                    : make_cont((lut2_word - (base16 - kContBase16)) / 4u);
    for (uint32_t i = first_long_prefix; i < (uint32_t)kLut1Size; ++i) lut1[i] = cont;

    for (int i = 0; i < nsym; ++i) {
        const int l = len_of[i];
        const uint32_t e = make_entry(spec.vals[i], l, is_dc);
        if (l <= kLutBits) {
            const uint32_t first = (uint32_t)code_of[i] << (kLutBits - l);
            const uint32_t n = 1u << (kLutBits - l);
            for (uint32_t j = 0; j < n; ++j)
                if (first + j < first_long_prefix) lut1[first + j] = e;
        } else {
            const uint32_t first = ((uint32_t)code_of[i] << (16 - l)) - base16;
            const uint32_t n = 1u << (16 - l);
            for (uint32_t j = 0; j < n; ++j)
                if (first + j < lut2_size) lut2[first + j] = e - consumed;
        }
    }
    // Pair fields (AC tables): if the index bits behind a complete first symbol hold another
    // complete symbol of the same table, record its length and advance.  The first symbol must
    // not be an end-of-block (the next symbol would be a DC symbol of another table).
    if (!is_dc) {
        // singles first: a pair entry's own fields hold totals, so the second symbol is looked up
        // in a copy of the table that has not been paired yet
        std::vector<uint32_t> single(lut1, lut1 + first_long_prefix);
        for (uint32_t i = 0; i < first_long_prefix; ++i) {
            const uint32_t e1 = single[i];
            if (e1 < kEntryMarkerLimit) continue;
            const uint32_t s1 = entry_skip(e1);
            if (s1 >= (uint32_t)kLutBits || entry_adv(e1) == 64u) continue;
            const uint32_t rem = (uint32_t)kLutBits - s1;
            const uint32_t i2 = (i << s1) & ((uint32_t)kLut1Size - 1u);
            if (i2 >= first_long_prefix) continue;
            const uint32_t e2 = single[i2];
            if (e2 < kEntryMarkerLimit || entry_skip(e2) > rem) continue;
            lut1[i] = with_pair(e1, entry_skip(e2), entry_adv(e2));
        }
        // singles half: what the write pass decodes with, and the state-only pass where a pair
        // could straddle the end of its span or of a data unit
        uint32_t* singles = lut1 + kLut1Size;
        for (uint32_t i = 0; i < (uint32_t)kLut1Size; ++i) singles[i] = i < first_long_prefix ? single[i] : lut1[i];
    }
    // short codes that fall inside the long region's prefixes cannot exist (canonical order puts
    // every short code numerically below every long code), so the two levels never conflict.
    return (uint32_t)(header_at + kLutHeader);
}

} // namespace jb2
#endif
