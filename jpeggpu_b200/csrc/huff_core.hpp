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
//  * the tables of the current data unit and of the next one come from one 16-byte
//    shared-memory record per MCU slot, fetched in the shadow of the LUT probe; tables, slots
//    and words are addressed through policy-defined handles (shared-memory byte addresses on
//    the device), so no index-to-address arithmetic sits in the loop.
//
// Memory access goes through a policy type `Mem` (shared memory on the device, plain arrays on
// the host):
//   lut(tbl, i)    -> entry i of the table with handle tbl;  single_of(tbl) -> its singles half
//   cont_table(e)  -> handle of the table a continuation entry points to (long codes)
//   slot_handle(c) -> handle of MCU slot c;  rec(h) -> SlotRec of that slot
//   slot_tbls(c)   -> dc table slot | ac table slot << 8   (slow path only)
//   slow(slot)     -> HuffSlow
//   put(slot, v)   -> store a non-zero coefficient (slot = data unit * 64 + zig-zag index,
//                     segment relative; slot % 64 == 0 is a DC difference)
//   word_handle(i) -> handle of big-endian word i of the scan window the policy was built for;
//   word(h)        -> that word; consecutive words are kWordStep handles apart
//   probe(tbl, peek) -> primary-LUT entry for the 32 stream bits `peek` (index = its top kLutBits)
//   slide_peek(w0, w1, w2, wn, pp, pn) -> move the window from bit position pp to pn (at most one
//                     word further) and return the 32 stream bits at pn
//   step<WRITE>(...) -> the bookkeeping of one symbol (generic_step, or the policy's own form)
#ifndef JB2_HUFF_CORE_HPP_
#define JB2_HUFF_CORE_HPP_

#include "common.hpp"

namespace jb2 {

// Per MCU slot c: the tables of its data unit and what follows it.  The fields are HANDLES whose
// meaning belongs to the memory policy (host: word offsets / indices; device: shared-memory byte
// addresses), so that the symbol loop never converts an index into an address.
struct SlotRec {
    uint32_t x; // ac table of slot c
    uint32_t y; // dc table of slot next(c)
    uint32_t z; // handle of slot next(c)
    uint32_t w; // dc table of slot c (low 24 bits) | c << 24
};

// canonical search for codes the two LUT levels do not cover (pathological tables only)
JB2_HD uint32_t slow_entry(const HuffSlow& t, uint32_t peek16)
{
    int l = kLutBits + 1;
    for (; l < 16; ++l)
        if (peek16 < t.limit[l]) break;
    const uint32_t code = peek16 >> (16 - l);
    const uint8_t sym = t.huffval[(uint8_t)(t.valoff[l] + (int)code)];
    return make_entry(sym, l, t.is_dc != 0);
}

JB2_HD int extend_value(uint32_t bits, uint32_t size)
{
    // T.81 F.2.2.1 (src/decode_huffman.cu:196-200); size >= 1
    return bits < (1u << (size - 1)) ? (int)bits - (int)(1u << size) + 1 : (int)bits;
}

// WRITE pass, portable form: the coefficient of the symbol whose entry is `e` (window `peek`, slot
// counter `qq` before the symbol) goes to the policy's put().  DC and AC alike: a zero value needs
// no store (every sink is pre-zeroed), the coefficient keeps its zig-zag index (the IDCT
// un-zig-zags in registers); continuations and kEntrySlow have size 0.
template <class Mem>
JB2_HD void put_symbol_generic(const Mem& M, uint32_t e, uint32_t peek, uint32_t qq)
{
    const uint32_t size = entry_size(e);
    const uint32_t pos = qq + entry_adv(e) - 1u;
    if (size != 0 && ((pos ^ qq) & 64u) == 0) {
        const uint32_t bits = (peek << (entry_skip(e) - size)) >> (32 - size);
        M.put(pos, extend_value(bits, size));
    }
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

// Window slide: when the bit position crosses a word boundary the window moves one word on.
// The window holds the two words being decoded (w0, w1) and the one after them (w2), which is
// loaded at slide time for the NEXT slide - at least 32 bits of symbols later - so the load
// latency never sits on the symbol-to-symbol dependency chain.  Policies provide
// slide_peek(w0, w1, w2, wn, pp, pn); this is the portable form.
template <class Mem>
JB2_HD void slide_window_generic(const Mem& M, uint32_t& w0, uint32_t& w1, uint32_t& w2,
                                 uint32_t& wn, uint32_t pp, uint32_t pn)
{
    if (((pn ^ pp) & 32u) != 0) {
        w0 = w1;
        w1 = w2;
        w2 = M.word(wn);
        wn += Mem::kWordStep;
    }
}

// Decode the symbols that start in [p, end).  State: p bit position inside the policy's word
// window, q linear slot counter (q & 63 = zig-zag index), c MCU slot of the current data unit.
// WRITE: hand every non-zero coefficient (DC differences included) to the policy's put(slot, v),
// slot = q of the coefficient (data unit q >> 6, zig-zag index q & 63), while q < q_limit.
//
// One step: position += skip(e), slot counter += advance(e), end-of-data-unit bookkeeping (EOB /
// slot 63: the counter snaps to the boundary, the MCU slot and the table move on), probe the table
// for the symbol that follows.  The loop's ONLY branch is its back edge, and its condition - the
// position moved and is still below `end` - is known at the start of the step, long before the
// branch issues (a predicate that comes fresh out of the table probe costs ~20 cycles of
// predicate-to-branch latency per symbol on sm_100).  What would need a branch on the entry is
// data instead:
//  * a code longer than kLutBits takes two ordinary steps: its primary entry is a continuation
//    (common.hpp) that consumes the code's leading one bits and names the table of the next
//    probe - one more select on the table address.  (An exit from the loop instead makes the
//    lane wait for the slowest lane of its warp, which then waits for it: measured +50 % on the
//    write pass, +100 % on the sync pass of a photo with 0.6 % long codes.)  The span cannot end
//    between the two steps: the loop continues after a continuation whatever the position.
//    Only pathological tables (long-code region > kLut2Max) still leave the loop: their entry
//    kEntrySlow is a no-op step, noticed because the position did not move;
//  * the state-only pass (huff_sync; !WRITE) takes both symbols of a pair entry in one step,
//    which equals two single steps iff the first symbol stays inside its data unit and the
//    second starts before `end`.  Both hold by construction while the zig-zag index is < 48 (a
//    non-terminal AC symbol advances by at most 16) and more than kLutBits - 1 bits remain (the
//    first symbol of a pair is shorter than kLutBits).  Elsewhere the NEXT probe simply goes to
//    the singles half of the AC table: a select on the table address, nothing on the chain of
//    the bit position.  The write pass only ever sees singles (its slot records point there).
// The bookkeeping of one step in portable form (policies may provide an equivalent `step`):
// consumes entry `e` at position pp; returns bit 0 = the loop goes on, bit 1 = the position moved.
template <bool WRITE, class Mem>
JB2_HD uint32_t generic_step(const Mem& M, uint32_t e, uint32_t& pp, uint32_t& qq, uint32_t& ch,
                             uint32_t& tbl, const SlotRec& r, uint32_t& w0, uint32_t& w1,
                             uint32_t& w2, uint32_t& wn, uint32_t& peek, uint32_t end,
                             uint32_t end_fast, uint32_t q_limit)
{
    const uint32_t pn = pp + entry_skip(e);
    const uint32_t qn = qq + entry_adv(e);
    peek = M.slide_peek(w0, w1, w2, wn, pp, pn);
    const bool moved = pn != pp;
    pp = pn;
    // advance <= 64 (single) or <= 80 from an index < 48 (pair): one boundary at most
    const bool unit_done = ((qn ^ qq) & 64u) != 0;
    const bool pairs_ok = WRITE || (pn < end_fast && ((~qn) & 48u) != 0u); // index < 48
    const bool cont = (e & kEntryCont) != 0;
    tbl = cont ? M.cont_table(e) : unit_done ? r.y : pairs_ok ? r.x : M.single_of(r.x);
    qq = unit_done ? (qn & ~63u) : qn;
    ch = unit_done ? r.z : ch;
    const bool go_on = moved && (cont || (pn < end && (!WRITE || qq < q_limit)));
    return (go_on ? 1u : 0u) | (moved ? 2u : 0u);
}

template <bool WRITE, class Mem>
JB2_HD void decode_span(const Mem& M, uint32_t& p, uint32_t end, uint32_t& q, uint32_t& c,
                        uint32_t q_limit)
{
    if (p >= end || (WRITE && q >= q_limit)) return;
    uint32_t pp = p, qq = q;
    uint32_t ch = M.slot_handle(c);
    SlotRec r = M.rec(ch);
    const uint32_t end_fast = end >= (uint32_t)kLutBits ? end - (uint32_t)(kLutBits - 1) : 0u;
    uint32_t tbl = (qq & 63u) == 0 ? (r.w & 0xffffffu)
                   : (!WRITE && (pp >= end_fast || (qq & 48u) == 48u)) ? M.single_of(r.x) : r.x;
    uint32_t wn = M.word_handle((pp >> 5) + 3u);
    uint32_t w0 = M.word(wn - 3u * Mem::kWordStep), w1 = M.word(wn - 2u * Mem::kWordStep);
    uint32_t w2 = M.word(wn - Mem::kWordStep);
    uint32_t peek = funnel_left_wrap(w0, w1, pp);
    uint32_t e = M.probe(tbl, peek);
    for (;;) {
        uint32_t flags;
        do {
            if (WRITE) M.put_symbol(e, peek, qq);
            flags = M.template step<WRITE>(e, pp, qq, ch, tbl, r, w0, w1, w2, wn, peek, end, end_fast, q_limit);
            e = M.probe(tbl, peek); // entry of the next symbol
            r = M.rec(ch);
        } while (flags & 1u);
        if (flags & 2u) break;
        // the position did not move: kEntrySlow, a long code of a pathological table
        M.note_long_code();
        const uint32_t tbls = M.slot_tbls(r.w >> 24);
        e = slow_entry(M.slow((qq & 63u) == 0 ? (tbls & 0xffu) : (tbls >> 8)), peek >> 16);
    }
    p = pp;
    q = qq;
    c = r.w >> 24;
}

// Fills the per-slot records with host-style handles (word offsets and slot indices).
// singles: the AC handle points at the singles half of the table (write pass).
JB2_HD void make_slot_recs(const ScanGeom& g, SlotRec* recs, uint32_t* tbls, bool singles = false)
{
    for (int s = 0; s < kMaxDuPerMcu; ++s) {
        const int cur = s < g.dus_per_mcu ? s : 0;
        const int nxt = cur + 1 == g.dus_per_mcu ? 0 : cur + 1;
        recs[s].x = g.table_off[g.slot_ac[cur]] + (singles ? (uint32_t)kLut1Size : 0u);
        recs[s].y = g.table_off[g.slot_dc[nxt]];
        recs[s].z = (uint32_t)nxt;
        recs[s].w = g.table_off[g.slot_dc[cur]] | ((uint32_t)cur << 24);
        tbls[s] = (uint32_t)g.slot_dc[cur] | ((uint32_t)g.slot_ac[cur] << 8);
    }
}

// Plain-array policy (host emulation; also documents what the device policy must provide).
struct HostMem {
    const uint32_t* lut_;
    const SlotRec* recs_;
    const uint32_t* tbls_; // per MCU slot: dc table slot | ac table slot << 8
    const HuffSlow* slow_;
    const uint32_t* words_; // word 0 = first word of the restart segment
    const uint32_t* table_off_ = nullptr; // ScanGeom::table_off (long codes only)
    static constexpr uint32_t kWordStep = 1;
    JB2_HD uint32_t lut(uint32_t tbl, uint32_t i) const { return lut_[(uint32_t)(tbl + i)]; }
    JB2_HD uint32_t single_of(uint32_t tbl) const { return tbl + (uint32_t)kLut1Size; }
    JB2_HD uint32_t cont_table(uint32_t e) const { return entry_cont_vec16(e) * 4u; }
    JB2_HD uint32_t slot_handle(uint32_t c) const { return c; }
    JB2_HD SlotRec rec(uint32_t h) const { return recs_[h]; }
    JB2_HD uint32_t slot_tbls(uint32_t c) const { return tbls_[c]; }
    JB2_HD const HuffSlow& slow(uint32_t s) const { return slow_[s]; }
    JB2_HD uint32_t word_handle(uint32_t i) const { return i; }
    JB2_HD uint32_t word(uint32_t h) const { return words_[h]; }
    JB2_HD uint32_t probe(uint32_t tbl, uint32_t peek) const { return lut(tbl, peek >> (32 - kLutBits)); }
    JB2_HD uint32_t slide_peek(uint32_t& w0, uint32_t& w1, uint32_t& w2, uint32_t& wn, uint32_t pp, uint32_t pn) const
    {
        slide_window_generic(*this, w0, w1, w2, wn, pp, pn);
        return funnel_left_wrap(w0, w1, pn);
    }
    template <bool WRITE>
    JB2_HD uint32_t step(uint32_t e, uint32_t& pp, uint32_t& qq, uint32_t& ch, uint32_t& tbl,
                         const SlotRec& r, uint32_t& w0, uint32_t& w1, uint32_t& w2, uint32_t& wn,
                         uint32_t& peek, uint32_t end, uint32_t end_fast, uint32_t q_limit) const
    {
        return generic_step<WRITE>(*this, e, pp, qq, ch, tbl, r, w0, w1, w2, wn, peek, end, end_fast, q_limit);
    }
    int16_t* coef_ = nullptr; // WRITE: first data unit of the restart segment
    int16_t* dc_ = nullptr;
    JB2_HD void put_symbol(uint32_t e, uint32_t peek, uint32_t qq) const { put_symbol_generic(*this, e, peek, qq); }
    JB2_HD void put(uint32_t slot, int v) const
    {
        if ((slot & 63u) == 0) dc_[slot >> 6] = (int16_t)v;
        else coef_[slot] = (int16_t)v;
    }
    long* long_codes_ = nullptr; // optional statistics
    JB2_HD void note_long_code() const
    {
        if (long_codes_) ++*long_codes_;
    }
};

} // namespace jb2
#endif
