// common.hpp — types shared by the host planner and the sm_100a kernels.
//
// Vocabulary (reference glossary: src/decoder.cpp:46-63):
//   scan        one SOS worth of entropy-coded data
//   segment     the part of a scan between two restart markers (independent bit stream)
//   subsequence kSubBytes destuffed bytes handled by one thread; never straddles a segment
//   data unit   one 8x8 block of 64 coefficients (DU)
//   slot        one coefficient position in coded order: slot = du * 64 + zigzag index
#ifndef JB2_COMMON_HPP_
#define JB2_COMMON_HPP_

#include <stddef.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define JB2_HD __host__ __device__ __forceinline__
#else
#define JB2_HD inline
#endif

#if defined(__GNUC__) || defined(__CUDACC__)
#define JB2_UNLIKELY(x) __builtin_expect(!!(x), 0)
#else
#define JB2_UNLIKELY(x) (x)
#endif

namespace jb2 {

constexpr int kMaxComp = 4;       // JPEGGPU_MAX_COMP (include/jpeggpu/jpeggpu.h:33)
constexpr int kMaxScans = 4;      // one component may be coded once (src/defs.hpp:84)
constexpr int kMaxDuPerMcu = 10;  // T.81 B.2.3
constexpr int kMaxTables = 8;     // 4 DC + 4 AC (src/defs.hpp:91)

// Subsequence size. The reference fixes 128 B (src/decoder_defs.hpp:32); we default to 64 B so a
// 12 MP scan exposes 2x the threads.  Must be a multiple of 4.
#ifndef JB2_SUB_BYTES
#define JB2_SUB_BYTES 64
#endif
constexpr int kSubBytes = JB2_SUB_BYTES;
constexpr int kSubBits = kSubBytes * 8;
constexpr int kSubWords = kSubBytes / 4;
// Overlap: the first kOverlapSubs lanes of a sync warp only warm up the decoder state for the
// warp's remaining 32 - kOverlapSubs subsequences (kernels_huffman.cu).
#ifndef JB2_OVERLAP_SUBS
#define JB2_OVERLAP_SUBS 6
#endif
constexpr int kOverlapSubs = JB2_OVERLAP_SUBS;
constexpr int kChunkSubs = 32 - kOverlapSubs; // subsequences whose records one sync warp owns

// Destuff tiling.
constexpr int kDestuffThreads = 256;
constexpr int kDestuffBytesPerThread = 16;
constexpr int kDestuffTile = kDestuffThreads * kDestuffBytesPerThread; // stuffed bytes per CTA

// ---- Huffman lookup tables -------------------------------------------------------------------
// One table = header (kLutHeader words: [0] base16, [1] secondary size, [2] word offset of the
// secondary LUT from the primary one, [3] 1 if a singles half follows the primary LUT)
//   + primary LUT (2^kLutBits words, indexed by the next kLutBits bits; AC tables: pair entries)
//   + AC tables only: the same LUT with every entry reduced to its first symbol ("singles half")
//   + secondary LUT (indexed by peek16 - base16) for codes longer than kLutBits
//   + slow-path arrays (HuffSlow) for pathological tables whose long-code region exceeds kLut2Max.
// Entry (32 bit): bits 31-27 total bits to skip (code length + magnitude bits, >= 1),
//                 bits 11-8 magnitude size, bits 6-0 slot advance (run + 1; 64 = end of block).
// The skip field sits at the top so that "position += skip" is a single shift-add and the
// advance at the bottom so that it is a single mask.
// A code longer than kLutBits has no branch of its own in the symbol loop; it takes two steps:
// the primary entry is a CONTINUATION (kEntryCont set): skip = kContBits - every long code starts
// with that many one bits, because the long-code region is the top <= kLut2Max values of the
// 16-bit code space - advance 0, size 0, and bits 26-12 = where the next probe goes instead of
// the regular table (in 16-byte units from the start of the LUT pool): the base of the table's
// secondary LUT, indexed like a primary one by the kLutBits bits that follow.  (The base lies
// before the entries that can actually be reached: index i stands for the 16-bit prefix
// 0xF800 + i, and prefixes below base16 never arrive there.)  Secondary entries are ordinary
// ones whose skip no longer counts the kContBits bits already consumed.
// kEntrySlow (all other fields zero; pathological tables whose long-code region exceeds kLut2Max)
// makes the step a no-op: the symbol loop notices that the position did not move and leaves.
#ifndef JB2_LUT_BITS
#define JB2_LUT_BITS 11
#endif
constexpr int kLutBits = JB2_LUT_BITS;
constexpr int kLut1Size = 1 << kLutBits;
constexpr int kLut2Max = kLut1Size; // a secondary LUT is probed like a primary one
constexpr int kLutHeader = 8;
constexpr uint32_t kContBits = 16 - kLutBits;           // 5: one bits consumed by a continuation entry
constexpr uint32_t kContBase16 = 0x10000u - kLut2Max;   // 0xF800: first prefix a secondary LUT can hold
constexpr uint32_t kEntryCont = 1u << 7;                // flag: continuation entry
constexpr uint32_t kEntrySlow = 2u << 12;               // canonical length search (no-op step)
constexpr uint32_t kEntryMarkerLimit = 1u << 27; // e < limit  <=>  marker

struct HuffSlow {       // canonical decode data (T.81 F.2.2.3), used by the slow path only
    uint32_t limit[17]; // limit[l] = (maxcode[l] + 1) << (16 - l), l = 1..16; 0 if no codes
    int32_t valoff[17]; // huffval index of first code of length l, minus its code
    uint8_t huffval[256];
    uint8_t is_dc;
    uint8_t pad[3];
};

JB2_HD uint32_t make_entry(unsigned symbol, int code_len, bool is_dc)
{
    unsigned size, adv;
    if (is_dc) {
        size = symbol & 15u; // valid streams: 0..11
        adv = 1;
    } else {
        const unsigned run = symbol >> 4;
        size = symbol & 15u;
        // src/decode_huffman.cu:236-242: size 0 is ZRL when run == 15, otherwise end of block
        adv = size ? run + 1 : (run == 15 ? 16u : 64u);
    }
    const unsigned skip = (unsigned)code_len + size; // <= 16 + 15
    return (skip << 27) | (size << 8) | adv;
}
JB2_HD uint32_t make_cont(uint32_t pool_off_vec16) // pool_off_vec16: secondary base / 16 bytes
{
    return (kContBits << 27) | (pool_off_vec16 << 12) | kEntryCont;
}
JB2_HD uint32_t entry_cont_vec16(uint32_t e) { return (e >> 12) & 0x7fffu; }
JB2_HD uint32_t entry_skip(uint32_t e) { return e >> 27; }
JB2_HD uint32_t entry_adv(uint32_t e) { return e & 127u; }
JB2_HD uint32_t entry_size(uint32_t e) { return (e >> 8) & 15u; }
// Pair entries (primary AC tables): when the LUT index also holds a complete second AC symbol
// behind the first, bits 26..22 = the second symbol's length in bits (code + magnitude), bits
// 21..15 = its slot advance, and the skip / advance fields hold the TOTALS of both symbols (the
// size field stays that of the first).  The state-only pass (huff_sync) consumes both symbols
// with one probe and without looking at the pair fields; the write pass first reduces an entry
// to its first symbol (huff_core.hpp first_symbol, or once per table at staging time).
JB2_HD uint32_t entry_skip2(uint32_t e) { return (e >> 22) & 31u; }
JB2_HD uint32_t entry_adv2(uint32_t e) { return (e >> 15) & 127u; }
JB2_HD uint32_t with_pair(uint32_t e, uint32_t skip2, uint32_t adv2)
{
    return (e | (skip2 << 22) | (adv2 << 15)) + (skip2 << 27) + adv2;
}

// ---- device-visible descriptors ---------------------------------------------------------------

struct SegDesc {          // one per restart segment, built by the host walk
    uint32_t sub_off;     // first subsequence of the segment
    uint32_t nbytes;      // destuffed data bytes in the segment
    uint32_t kept_before; // destuffed data bytes in all earlier segments of the scan
    uint32_t src_begin;   // stuffed offset (relative to scan start) of the first byte
};

// Threads that share one subsequence in the write pass; the sync pass records the decoder
// state at the kWriteParts - 1 inner boundaries ("mid states").
#ifndef JB2_WRITE_PARTS
#define JB2_WRITE_PARTS 2
#endif
constexpr int kWriteParts = JB2_WRITE_PARTS;
static_assert(kWriteParts == 2 || kWriteParts == 4, "two or four threads per subsequence");
constexpr int kRecWords = kWriteParts == 2 ? 2 : 4; // 64-bit words per sync record

struct SubEntry { // per-subsequence sync record: kRecWords 8-byte words
    uint16_t g;   // packed entry state (bit 15: first subsequence of a restart segment)
    uint16_t x;   // packed exit state given g
    uint32_t n;   // slots advanced inside the subsequence given g
    // mid[k]: packed state at inner boundary k | slots advanced before it << 16; two per word
    uint32_t mid[2 * (kRecWords - 1)];
};
struct Mids {
    uint32_t v[kWriteParts - 1];
};
constexpr uint16_t kStateHead = 0x8000u;
constexpr uint16_t kStateMask = 0x7fffu;

// Totals of a scan, written by the destuff tile that holds the end of the scan and read by every
// later kernel: the host plans grids and buffers from upper bounds and need not know them.
struct ScanInfo {
    uint32_t num_subs;     // N
    uint32_t num_segments; // restart segments (the last one closed by the end of the scan)
    uint32_t scan_end;     // stuffed offset of the byte that ended the scan (marker code or end of data)
    uint32_t status;       // kScan* flags
};
constexpr uint32_t kScanUnterminated = 1u; // the data ran out before a terminating marker
constexpr uint32_t kScanOverflow = 2u;     // more segments / subsequences than planned (damaged stream)

struct TileDesc {         // one per kDestuffTile stuffed bytes, built by the host walk (tests only)
    uint32_t kept_before; // destuffed data bytes of the scan before the tile's first byte
    uint32_t segs_before; // restart markers completed before the tile's first byte
};

// packed state: overflow bits past the boundary (5) | DU index in MCU (4) | zig-zag index (6)
JB2_HD uint16_t pack_state(unsigned o, unsigned c, unsigned z)
{
    return (uint16_t)(o | (c << 5) | (z << 9));
}

struct ScanGeom { // by-value kernel parameter describing one scan
    int num_comp;         // components in the scan
    int dus_per_mcu;      // K
    int mcus_x, mcus_y;
    int total_mcus;
    int restart_interval; // MCUs per segment; == total_mcus when no DRI
    int num_segments;
    uint32_t num_subs;    // N
    uint32_t scan_bytes;  // stuffed bytes Sb
    uint32_t scan_avail;  // bytes handed to the device from the start of the scan: Sb + the two
                          // bytes of the marker that ends it (host walk), or everything up to
                          // the end of the file (device walk)
    int num_tables;
    // per DU slot of the MCU
    uint8_t slot_dc[kMaxDuPerMcu]; // table slot of the DC table
    uint8_t slot_ac[kMaxDuPerMcu];
    uint8_t slot_comp[kMaxDuPerMcu]; // scan component of the slot
    // per scan component
    int frame_comp[kMaxComp];
    int ssx[kMaxComp], ssy[kMaxComp]; // DUs per MCU in x / y (1 for non-interleaved)
    int du_off[kMaxComp];             // first slot of the component inside the MCU
    int blocks_x[kMaxComp], blocks_y[kMaxComp]; // coded plane size in DUs (MCU padded)
    int size_x[kMaxComp], size_y[kMaxComp];     // output plane size in samples
    uint32_t du_prefix[kMaxComp + 1];           // prefix sum of blocks_x*blocks_y
    uint32_t table_off[kMaxTables];             // word offset of each table's primary LUT in the pool
    uint32_t lut_pool_words;                    // total 32-bit words in the pool
    uint32_t lut_compact_words;                 // words huff_write stages: header + one LUT half + secondary per table
};

struct OutPlanes {
    uint8_t* ptr[kMaxComp]; // indexed by scan component
    int pitch[kMaxComp];
    uint8_t q[kMaxComp][64]; // natural-order quantisation table per scan component
};

// Interleaved RGB destination of the fused IDCT + colour conversion (jpeggpu_b200_decode_rgbi).
struct RgbOut {
    uint8_t* ptr;  // 3 bytes per pixel, rows `pitch` bytes apart
    int pitch;
    int width, height;
    int cy, cb, cr; // scan component that carries Y / Cb / Cr; cb < 0: single grey component
    uint8_t q[kMaxComp][64]; // natural-order quantisation table per scan component
};

// Planes that already exist (jpeggpu_b200_convert_rgbi): frame component order.
struct PlanarIn {
    const uint8_t* ptr[3];
    int pitch[3];
    int num_comp;  // 1 or 3
    int ssx, ssy;  // luma samples per chroma sample
    bool aligned;  // every plane pointer and pitch is a multiple of 8 (set by the launcher)
};

struct CoefPlanes {
    int16_t* ptr[kMaxComp]; // indexed by scan component
    int stride[kMaxComp];   // in int16 elements
};

// Diagnostic build (-DJB2_TIMELINE, tools/timeline.py): globaltimer stamps of every kernel's
// first start / last end in front of the control block (never cleared by the kernels).
#ifdef JB2_TIMELINE
constexpr int kTimelineWords = 32;
#else
constexpr int kTimelineWords = 0;
#endif

// Control block at the front of the per-scan work area (zeroed every decode).
struct Control {
#ifdef JB2_TIMELINE
    uint32_t tl[kTimelineWords];
#endif
    uint32_t relinked;    // warp-boundary mismatches healed inside huff_sync
    uint32_t slow_repairs; // records rewritten by the sequential repair in huff_sync's last CTA
    uint32_t link_rounds;
    uint32_t link_verify_ns; // tail of the last huff_sync CTA: verification (+ repair) (diagnostics)
    uint32_t need_verify;    // a chunk's exit state changed after its right neighbour may have read it
    uint32_t write_stage_ns;  // slowest CTA of huff_write: staging / decode / flush (diagnostics)
    uint32_t write_decode_ns;
    uint32_t write_flush_ns;
    uint32_t sync_stage_ns;   // slowest CTA of huff_sync: staging / fix-point iteration
    uint32_t sync_iter_ns;
    uint32_t reserved1[2];
    uint32_t dc_ticket;       // dc_scan: tiles are numbered in the order their CTAs start
    uint32_t write_ticket;    // huff_write: likewise
    uint32_t sync_done;       // huff_sync CTAs that have finished (the last one verifies)
    uint32_t pad[1];
};

// zig-zag -> natural order (T.81 Figure A.6; same table as src/defs.hpp:94-102)
#define JB2_NATURAL_ORDER                                                                      \
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, \
        20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44,  \
        51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63

} // namespace jb2
#endif
