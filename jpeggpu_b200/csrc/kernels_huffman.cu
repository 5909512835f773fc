// kernels_huffman.cu — self-synchronising parallel Huffman decode (reference stages a4-a11,
// src/decode_huffman.cu:413-880; algorithm family of arXiv 2111.09219).
//
// The reference runs "flow into the next subsequence until your state equals the stored one" per
// 256-thread block with a block-wide reduction and two barriers per round, then a second kernel
// across blocks, a CUB scan-by-key, a copy kernel and the write kernel.  Here:
//
//   huff_sync   one WARP owns kChunkSubs consecutive subsequences and additionally decodes the
//               kOverlapSubs before them as warm-up.  The CTA stages its slice of the destuffed
//               scan and the decode LUTs in shared memory.  Every lane decodes its subsequence
//               from the cold state, then the warp iterates "take the exit state of the lane on
//               your left (one shuffle), re-decode if your entry state changed" until no lane
//               changes (one vote per round, no barrier).  After r rounds lane i holds the state a
//               decoder cold-started r subsequences earlier would have; the warm-up lanes absorb
//               the cold start of lane 0.  The warp then publishes the exit state of its last
//               subsequence in a one-word mailbox, reads its left neighbour's mailbox, and if that
//               disagrees with the entry state it had assumed it re-runs the iteration from the
//               corrected state (rare).  A lane keeps the evaluation it held before the current
//               one (one-entry memo), so flipping back to an earlier entry state costs nothing.
//               One 16-byte record {entry, exit, slots, mid state} per subsequence.
//               An extra warp per CTA zero-fills the coefficient buffer in the prologue (buffers
//               that fit L2; larger ones are cleared by huff_write, stretch by stretch).
//               Tables and scan window arrive by cp.async.bulk + mbarrier.
//               Only when a chunk re-published a different exit state (or a mailbox wait timed
//               out) does the CTA that finishes last verify every chunk boundary (exit of the left
//               record == entry of the right one) and repair sequentially from each break;
//               otherwise every record is exact by induction from the segment starts.
//   huff_write  decode from the verified entry states, kWriteParts (2) threads per subsequence (the record
//               holds the state at its middle).  The first output slot of every subsequence is
//               a segmented prefix sum of the records' slot counts, computed inside this kernel
//               with a look-back over per-CTA aggregates (replaces the reference's CUB
//               ExclusiveScanByKey + assign_sinfo_n, src/decode_huffman.cu:827-869).  Non-zero
//               coefficients are collected in bounded per-thread lists in shared memory (what
//               does not fit goes straight to global memory) and scattered afterwards into the
//               zero-filled buffer, zig-zag index kept (the IDCT un-zig-zags
//               in registers); DC differences go to a compact per-DU array
//               (src/decode_huffman.cu:627-682).
#include "huff_threads.hpp"
#include "kernels.hpp"

namespace jb2 {
namespace {

// huff_sync CTAs: kHuffThreads / 32 warps for one image (many small CTAs start decoding sooner and
// spread over all SMs), kSyncWarpsBatch warps for batched launches (the tables are staged once per
// CTA: with thousands of CTAs per launch that staging is what the SMs spend their time on; 1 GPU,
// 4096 files of 2 MP: 70 k -> 82 k images/s).
constexpr int kSyncWarps = kHuffThreads / 32;
#ifndef JB2_SYNC_WARPS_BATCH
#define JB2_SYNC_WARPS_BATCH 8
#endif
constexpr int kSyncWarpsBatch = JB2_SYNC_WARPS_BATCH;
// + one warp per CTA that zero-fills the CTA's slice of the coefficient buffer and of the DC array
// (huff_write scatters the non-zero values into them).  Measured on B200, 12 MP, bench.py
// ms_per_step: 0 = every thread fills in the prologue, in competition with the destuff kernel
// (both done at 15-16 us, destuff alone takes 11.7) 0.1293; 1 = the extra warp fills while the
// others iterate (they are latency bound, the store path is idle) 0.1270 / 0.1256 on a second
// box; 2 = the extra warp fills in the prologue, at a quarter of the store rate of mode 0, and is
// idle afterwards 0.1243 on that second box.
// When the coefficient buffer does not fit the L2 cache the writers clear their own stretches
// instead (ScanBufs::fill_in_write, huff_write_body) and the extra warp has nothing to do.
#ifndef JB2_FILL_WARP
#define JB2_FILL_WARP 2
#endif
constexpr int kFillWarps = (JB2_FILL_WARP == 1 || JB2_FILL_WARP == 2) ? 1 : 0;
constexpr int kSyncThreads = (kSyncWarps + kFillWarps) * 32;
constexpr int kSyncThreadsBatch = (kSyncWarpsBatch + kFillWarps) * 32;
// Diagnostic knobs (tests build variants): how long a chunk waits for its left neighbour's exit
// state, and whether the last CTA verifies the chunk boundaries even when no chunk asked for it.
#ifndef JB2_MAILBOX_SPINS
#define JB2_MAILBOX_SPINS (1 << 18)
#endif
#ifndef JB2_ALWAYS_VERIFY
#define JB2_ALWAYS_VERIFY 0
#endif
#ifndef JB2_WRITE_MIN_WARPS
#define JB2_WRITE_MIN_WARPS 4
#endif
constexpr int kMailboxSpins = JB2_MAILBOX_SPINS;
constexpr bool kAlwaysVerify = JB2_ALWAYS_VERIFY != 0;
constexpr int kWinMargin = 4;                                   // words past the last subsequence
__host__ __device__ constexpr int sync_useful(int warps) { return warps * kChunkSubs; } // records written per CTA
__host__ __device__ constexpr int sync_win_words(int warps)                             // words staged per CTA
{
    return (sync_useful(warps) + kOverlapSubs) * kSubWords + kWinMargin;
}
constexpr int kWriteThreads = 160 * kWriteParts; // at most; kWriteParts threads per subsequence
constexpr int kSubsPerWarp = 32 / kWriteParts;
// Non-zero coefficients a write thread keeps in its shared-memory list.  Half a subsequence can
// hold kSubBits / (2 * kWriteParts) + 2 of them in theory (2 bits each); real streams stay far
// below (tools/sim/list_stats.cpp: mean 42-46, maximum 57-61 per 32 bytes on the test images), and
// lists sized for the theoretical maximum cost 210 KB per CTA: one CTA per SM, four waves at
// 48 MP.  A list that fills up sends the rest of its coefficients straight to global memory from
// inside the symbol loop (fire-and-forget stores).  Measured on B200, huff_write for 12 MP /
// IMG_6510 / 48 MP / 2 MP in us: no list 35.5 / 69.1 / 146.9 / 30.5, 16 entries 36.6 / 64.8 /
// 144.7 / 31.1, 32 entries 37.1 / 50.8 / 131.2 / 32.6, 48 entries 38.4 / 50.3 / 134.3 / 34.5,
// 64 entries 45.4 / 67.0 / 149.5 / 30.7, 130 entries (never overflows) 37.5 / 58 / 152 / 34:
// the kernel sits on the scattered 2-byte stores whichever way they leave (~3 us per megapixel).
#ifndef JB2_WRITE_LIST
#define JB2_WRITE_LIST 32
#endif
constexpr int kListCap = JB2_WRITE_LIST < kSubBits / (2 * kWriteParts) + 2 ? JB2_WRITE_LIST
                                                                         : kSubBits / (2 * kWriteParts) + 2;
constexpr int kListStride = kListCap | 1;                          // words between the lanes' lists
constexpr int kWarpListBytes = (32 * kListStride * 4 + 15) & ~15;
constexpr int kWarpSmem = kWarpListBytes;                          // shared memory per write warp
static_assert(kSubBytes <= 128, "slot offsets inside half a subsequence must fit 16 bits");
static_assert(sync_win_words(kSyncWarps) % 4 == 0 && sync_win_words(kSyncWarpsBatch) % 4 == 0 && kSubWords % 4 == 0,
              "window must be uint4 sized");

// Bulk staging (Blackwell/Hopper copy engine): the decode tables of huff_sync and the scan window of
// both kernels are contiguous in global memory, so ONE cp.async.bulk each, completed on an mbarrier,
// replaces the per-thread 16-byte load/store loops (3456 + 3620 of them per huff_sync CTA); the
// table copy then overlaps the zero-fill of the coefficient buffer.  -DJB2_BULK_STAGE=0 restores
// the loops.
#ifndef JB2_BULK_STAGE
#define JB2_BULK_STAGE 1
#endif
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
// Bounded: a copy that never completes (it cannot, short of a hardware fault) must not hang the GPU;
// each try_wait already blocks for a hardware-defined time slice.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    for (int spin = 0; spin < (1 << 20); ++spin) {
        uint32_t done;
        asm volatile("{\n"
                     ".reg .pred p;\n"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
                     "selp.u32 %0, 1, 0, p;\n"
                     "}"
                     : "=r"(done)
                     : "r"(bar), "r"(parity)
                     : "memory");
        if (done) return;
    }
}

struct __align__(16) SharedTables {
    SlotRec recs[kMaxDuPerMcu]; // device handles: shared-memory byte addresses
    uint32_t tbls[kMaxDuPerMcu];
    uint32_t tbase[kMaxTables]; // shared address of every table's primary LUT (header in front)
};

extern __shared__ __align__(16) uint4 s_dyn[]; // [window words | LUT pool]

__device__ __forceinline__ unsigned long long now_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ uint32_t smem_addr(const void* p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr)
{
    uint32_t v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ SlotRec lds_rec(uint32_t addr)
{
    SlotRec r;
    asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
        : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
        : "r"(addr));
    return r;
}

// Converts the host-style slot records into shared-memory addresses (thread 0, before the
// barrier in stage_window()).  st.tbase[] must be set.  SINGLE: AC handles point at singles.
__device__ __forceinline__ void make_device_recs(const ScanGeom& g, SharedTables& st, bool staged_singles)
{
    const uint32_t recs_addr = smem_addr(st.recs);
    for (int s = 0; s < kMaxDuPerMcu; ++s) {
        const int cur = s < g.dus_per_mcu ? s : 0;
        const int nxt = cur + 1 == g.dus_per_mcu ? 0 : cur + 1;
        SlotRec& r = st.recs[s];
        // the compact staging of huff_write holds ONLY the singles half, at the table base
        r.x = st.tbase[g.slot_ac[cur]];
        r.y = st.tbase[g.slot_dc[nxt]];
        r.z = recs_addr + (uint32_t)nxt * (uint32_t)sizeof(SlotRec);
        r.w = st.tbase[g.slot_dc[cur]] | ((uint32_t)cur << 24);
        st.tbls[s] = (uint32_t)g.slot_dc[cur] | ((uint32_t)g.slot_ac[cur] << 8);
    }
    (void)staged_singles;
}

// Stage the decode tables and words [first_word, first_word + win_words) of the destuffed scan
// (zero outside [0, total_words)).
// Staging in two steps so that the first can run before the producer of the destuffed scan has
// finished (programmatic dependent launch): the decode tables come from the upload block...
// SINGLE (huff_write): per table only header + ONE primary LUT (the singles half of an AC table)
// + secondary LUT are staged, back to back, so the pool costs the write kernel no more shared
// memory than a plain table set; huff_sync takes the pool as it is.
template <bool SINGLE>
__device__ __forceinline__ void stage_tables(const ScanGeom& g, const ScanBufs& b, SharedTables& st,
                                             const int WIN_WORDS, const uint32_t bar_tables = 0)
{
    uint4* ldst = s_dyn + WIN_WORDS / 4;
    const uint32_t lut_addr = smem_addr(ldst);
    if (!SINGLE) {
        const uint32_t lut_vec = (g.lut_pool_words + 3) / 4; // pool is padded to 16 B by the planner
        const uint4* lsrc = reinterpret_cast<const uint4*>(b.lut_pool);
        if (JB2_BULK_STAGE) {
            // one bulk copy, waited for in stage_window (bar_tables: 0 = none in flight)
            if (threadIdx.x == 0) {
                mbar_init(bar_tables, 1);
                mbar_expect_tx(bar_tables, lut_vec * 16u);
                bulk_g2s(lut_addr, lsrc, lut_vec * 16u, bar_tables);
            }
        } else {
            for (uint32_t i = threadIdx.x; i < lut_vec; i += blockDim.x) ldst[i] = __ldg(lsrc + i);
        }
        if (threadIdx.x == 0) {
            for (int t = 0; t < g.num_tables; ++t) st.tbase[t] = lut_addr + g.table_off[t] * 4u;
            make_device_recs(g, st, false);
        }
        return;
    }
    uint32_t at = 0; // 16-byte vectors staged so far (uniform)
    for (int t = 0; t < g.num_tables; ++t) {
        const uint32_t* hdr = b.lut_pool + g.table_off[t] - kLutHeader;
        const uint4 h = __ldg(reinterpret_cast<const uint4*>(hdr)); // base16, n2, off2, has singles
        const uint32_t n2_vec = (h.y + 3u) / 4u;
        const uint4* lut1 = reinterpret_cast<const uint4*>(hdr + kLutHeader + (h.w ? kLut1Size : 0));
        const uint4* lut2 = reinterpret_cast<const uint4*>(hdr + kLutHeader + h.z);
        uint4* dst = ldst + at;
        if (threadIdx.x == 0) {
            dst[0] = make_uint4(h.x, h.y, (uint32_t)kLut1Size, 0u);
            dst[1] = make_uint4(0u, 0u, 0u, 0u);
            st.tbase[t] = lut_addr + (at + kLutHeader / 4) * 16u;
        }
        dst += kLutHeader / 4;
        // continuation entries name the secondary LUT by its place in the pool: re-base them
        const uint32_t pool_vec = (g.table_off[t] + h.z) / 4u;                 // secondary LUT in the pool
        const uint32_t here_vec = at + (uint32_t)(kLutHeader + kLut1Size) / 4; // ... and in this staging
        const uint32_t rebase = ((here_vec - pool_vec) & 0x7fffu) << 12;
        auto fix = [rebase](uint32_t v) {
            return (v & kEntryCont) ? ((v & ~(0x7fffu << 12)) | ((v + rebase) & (0x7fffu << 12))) : v;
        };
        for (uint32_t i = threadIdx.x; i < (uint32_t)kLut1Size / 4; i += blockDim.x) {
            uint4 v = __ldg(lut1 + i);
            v.x = fix(v.x), v.y = fix(v.y), v.z = fix(v.z), v.w = fix(v.w);
            dst[i] = v;
        }
        dst += kLut1Size / 4;
        for (uint32_t i = threadIdx.x; i < n2_vec; i += blockDim.x) dst[i] = __ldg(lut2 + i);
        at += (uint32_t)(kLutHeader + kLut1Size) / 4 + n2_vec;
    }
    // tbase[] was written by thread 0 itself: no barrier needed before it builds the records
    if (threadIdx.x == 0) make_device_recs(g, st, true);
}
// ...the window of the scan from the destuff kernel's output.  Ends with a barrier.
__device__ __forceinline__ void stage_window(const ScanBufs& b, uint32_t num_subs,
                                             long long first_word, const int WIN_WORDS,
                                             const uint32_t bar_window = 0, const uint32_t bar_tables = 0)
{
    uint32_t* win = reinterpret_cast<uint32_t*>(s_dyn);
    const long long total_words = (long long)num_subs * kSubWords;
    const uint4* src = reinterpret_cast<const uint4*>(b.destuffed);
    if (JB2_BULK_STAGE && bar_window) {
        // words [lo, hi) of the window exist in the scan (all multiples of 4): one bulk copy; the
        // rest (first and last CTA of a scan only) is zero
        long long lo_l = first_word < 0 ? -first_word : 0;
        if (lo_l > WIN_WORDS) lo_l = WIN_WORDS;
        long long hi_l = total_words - first_word < WIN_WORDS ? total_words - first_word : WIN_WORDS;
        if (hi_l < lo_l) hi_l = lo_l; // a window behind the end of the scan (grids are sized from bounds)
        const int lo = (int)lo_l, hi = (int)hi_l;
        if (threadIdx.x == 0) {
            mbar_init(bar_window, 1);
            if (hi > lo) {
                // the scan was written by the previous grid through the generic proxy
                asm volatile("fence.proxy.async;" ::: "memory");
                mbar_expect_tx(bar_window, (uint32_t)(hi - lo) * 4u);
                bulk_g2s(smem_addr(win + lo), b.destuffed + first_word + lo, (uint32_t)(hi - lo) * 4u, bar_window);
            }
        }
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        for (int i = threadIdx.x; i < lo / 4; i += blockDim.x) reinterpret_cast<uint4*>(win)[i] = z;
        for (int i = hi / 4 + threadIdx.x; i < WIN_WORDS / 4; i += blockDim.x) reinterpret_cast<uint4*>(win)[i] = z;
        __syncthreads(); // the barrier words are initialised
        if (hi > lo) mbar_wait(bar_window, 0);
        if (bar_tables) mbar_wait(bar_tables, 0);
        return;
    }
    for (int i = threadIdx.x; i < WIN_WORDS / 4; i += blockDim.x) {
        const long long w = first_word + 4ll * i;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (w >= 0 && w + 4 <= total_words) v = __ldg(src + (w >> 2));
        reinterpret_cast<uint4*>(win)[i] = v;
    }
    __syncthreads();
}

struct DevMem {
    const SharedTables* st;
    const HuffSlow* slow_;
    uint32_t recs_addr;
    uint32_t win_addr;
    uint32_t lut_addr; // shared address of the staged LUT pool
    static constexpr uint32_t kWordStep = 4;
    __device__ __forceinline__ uint32_t single_of(uint32_t tbl) const { return tbl + (uint32_t)kLut1Size * 4u; }
    // generic_step (huff_core.hpp) with every decision as a select: ptxas turns the C++ form into
    // three branches per symbol.  Only the loop's back edge branches, on a predicate that is
    // ready long before it issues.
    template <bool WRITE>
    __device__ __forceinline__ uint32_t step(uint32_t e, uint32_t& pp, uint32_t& qq, uint32_t& ch,
                                             uint32_t& tbl, const SlotRec& r, uint32_t& w0, uint32_t& w1,
                                             uint32_t& w2, uint32_t& wn, uint32_t& peek, uint32_t end,
                                             uint32_t end_fast, uint32_t q_limit) const
    {
        uint32_t flags;
        asm volatile(
            "{\n\t"
            ".reg .pred s, d, po, c, mv, go, ql;\n\t"
            ".reg .b32 t, a, b, pn, qn, x2, ct;\n\t"
            "shr.u32 t, %10, 27;\n\t"
            "add.u32 pn, %1, t;\n\t"
            "and.b32 t, %10, 127;\n\t"
            "add.u32 qn, %2, t;\n\t"
            // window: both candidates, then select; the registers slide off the chain
            "shf.l.wrap.b32 a, %6, %5, pn;\n\t"
            "shf.l.wrap.b32 b, %7, %6, pn;\n\t"
            "xor.b32 t, pn, %1;\n\t"
            "and.b32 t, t, 32;\n\t"
            "setp.ne.u32 s, t, 0;\n\t"
            "selp.b32 %9, b, a, s;\n\t"
            "@s mov.b32 %5, %6;\n\t"
            "@s mov.b32 %6, %7;\n\t"
            "@s ld.shared.u32 %7, [%8];\n\t"
            "@s add.u32 %8, %8, 4;\n\t"
            // end of the data unit?
            "xor.b32 t, qn, %2;\n\t"
            "and.b32 t, t, 64;\n\t"
            "setp.ne.u32 d, t, 0;\n\t"
            // pairs allowed for the next probe: zig-zag index < 48 and far enough from `end`
            "not.b32 t, qn;\n\t"
            "and.b32 t, t, 48;\n\t"
            "setp.ne.u32 po, t, 0;\n\t"
            "setp.lt.and.u32 po, pn, %15, po;\n\t"
            "and.b32 t, %10, 128;\n\t"
            "setp.ne.u32 c, t, 0;\n\t"
            // table of the next probe
            "add.u32 x2, %11, %17;\n\t"
            "selp.b32 t, %11, x2, po;\n\t"
            "selp.b32 t, %12, t, d;\n\t"
            "shr.u32 ct, %10, 8;\n\t"
            "and.b32 ct, ct, 0x7fff0;\n\t"
            "add.u32 ct, ct, %16;\n\t"
            "selp.b32 %4, ct, t, c;\n\t"
            // slot counter and MCU slot
            "and.b32 t, qn, 0xffffffc0;\n\t"
            "selp.b32 %2, t, qn, d;\n\t"
            "selp.b32 %3, %13, %3, d;\n\t"
            // loop control: moved && (cont || (pn < end && slot counter < q_limit))
            "setp.ne.u32 mv, pn, %1;\n\t"
            "setp.lt.u32 ql, %2, %18;\n\t"
            "setp.lt.and.u32 go, pn, %14, ql;\n\t"
            "or.pred go, go, c;\n\t"
            "and.pred go, go, mv;\n\t"
            "selp.u32 %0, 1, 0, go;\n\t"
            "@mv or.b32 %0, %0, 2;\n\t"
            "mov.b32 %1, pn;\n\t"
            "}"
            : "=&r"(flags), "+r"(pp), "+r"(qq), "+r"(ch), "=&r"(tbl), "+r"(w0), "+r"(w1), "+r"(w2), "+r"(wn),
              "=&r"(peek)
            : "r"(e), "r"(r.x), "r"(r.y), "r"(r.z), "r"(end), "r"(WRITE ? 0xffffffffu : end_fast),
              "r"(lut_addr), "r"(WRITE ? 0u : (uint32_t)kLut1Size * 4u), "r"(WRITE ? q_limit : 0xffffffffu));
        return flags;
    }
    __device__ __forceinline__ uint32_t slot_handle(uint32_t c) const
    {
        return recs_addr + c * (uint32_t)sizeof(SlotRec);
    }
    __device__ __forceinline__ SlotRec rec(uint32_t h) const { return lds_rec(h); }
    __device__ __forceinline__ uint32_t slot_tbls(uint32_t c) const { return st->tbls[c]; }
    __device__ __forceinline__ const HuffSlow& slow(uint32_t s) const { return slow_[s]; }
    __device__ __forceinline__ uint32_t word_handle(uint32_t i) const { return win_addr + i * 4u; }
    __device__ __forceinline__ uint32_t word(uint32_t h) const { return lds32(h); }
    // The probe address in two dependent instructions: mask the index bits, multiply-high-add.
    __device__ __forceinline__ uint32_t probe(uint32_t tbl, uint32_t peek) const
    {
        uint32_t addr;
        asm("mad.hi.u32 %0, %1, %2, %3;"
            : "=r"(addr)
            : "r"(peek & (0xffffffffu << (32 - kLutBits))), "r"(1u << (kLutBits + 2)), "r"(tbl));
        return lds32(addr);
    }
    __device__ __forceinline__ void note_long_code() const {}
    __device__ __forceinline__ void put_symbol(uint32_t, uint32_t, uint32_t) const {}
};

// Policy of huff_write: DevMem + the coefficient sink.  The symbol loop only appends
// {slot - slot0, value} to the thread's list in shared memory (one contiguous region per lane, an
// odd number of words apart); huff_write turns the lists into data units afterwards, all lanes
// working on one list at a time.
struct WriteMem : DevMem {
    mutable uint32_t list_ptr; // shared address of the next entry
    uint32_t list_limit;       // end of the list
    uint32_t slot0;            // first slot of this thread's half subsequence (segment relative)
    int16_t* coef_seg;         // overflow: coefficient buffer / DC array at the first data unit of
    int16_t* dc_seg;           // the restart segment
    // put_symbol_generic without a branch: value extraction, "is there a coefficient", "list or
    // overflow" and "DC or AC" are predicates of two predicated stores.  The C++ form costs two
    // data-dependent branches per symbol on predicates that come fresh from the table probe, and
    // the lanes of a warp take them differently (JB2_WRITE_PTX=0 restores it).
#ifndef JB2_WRITE_PTX
#define JB2_WRITE_PTX 1
#endif
#if !JB2_WRITE_PTX
    __device__ __forceinline__ void put(uint32_t slot, int v) const
    {
        if (list_ptr != list_limit) {
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(list_ptr),
                         "r"(((slot - slot0) << 16) | ((uint32_t)v & 0xffffu)));
            list_ptr += 4u;
        } else if ((slot & 63u) == 0) {
            dc_seg[slot >> 6] = (int16_t)v;
        } else {
            coef_seg[slot] = (int16_t)v;
        }
    }
#endif
    __device__ __forceinline__ void put_symbol(uint32_t e, uint32_t peek, uint32_t qq) const
    {
#if JB2_WRITE_PTX
        asm volatile(
            "{\n\t"
            ".reg .pred ps, pl, po, pd, pn;\n\t"
            ".reg .b32 sz, pos, t, u, bits, val, pk;\n\t"
            ".reg .b64 ad, bs;\n\t"
            "shr.u32 sz, %1, 8;\n\t"
            "and.b32 sz, sz, 15;\n\t"
            "and.b32 t, %1, 127;\n\t"
            "add.u32 pos, %3, t;\n\t"
            "sub.u32 pos, pos, 1;\n\t"
            // a coefficient: size != 0 and inside the data unit
            "xor.b32 t, pos, %3;\n\t"
            "and.b32 t, t, 64;\n\t"
            "setp.eq.u32 ps, t, 0;\n\t"
            "setp.ne.and.u32 ps, sz, 0, ps;\n\t"
            // magnitude bits (PTX shifts clamp: size 0 gives 0) and their sign extension
            "shr.u32 t, %1, 27;\n\t"
            "sub.u32 t, t, sz;\n\t"
            "shl.b32 bits, %2, t;\n\t"
            "sub.u32 t, 32, sz;\n\t"
            "shr.u32 bits, bits, t;\n\t"
            "sub.u32 t, sz, 1;\n\t"
            "shl.b32 u, 1, t;\n\t"
            "setp.lt.u32 pn, bits, u;\n\t"
            "shl.b32 t, 1, sz;\n\t"
            "sub.u32 t, bits, t;\n\t"
            "add.u32 t, t, 1;\n\t"
            "selp.b32 val, t, bits, pn;\n\t"
            // list entry, or straight to global memory when the list is full
            "setp.ne.and.u32 pl, %0, %4, ps;\n\t"
            "setp.eq.and.u32 po, %0, %4, ps;\n\t"
            "sub.u32 t, pos, %5;\n\t"
            "shl.b32 pk, t, 16;\n\t"
            "and.b32 t, val, 0xffff;\n\t"
            "or.b32 pk, pk, t;\n\t"
            "@pl st.shared.u32 [%0], pk;\n\t"
            "@pl add.u32 %0, %0, 4;\n\t"
            "and.b32 t, pos, 63;\n\t"
            "setp.eq.u32 pd, t, 0;\n\t"
            "shr.u32 t, pos, 6;\n\t"
            "selp.b32 t, t, pos, pd;\n\t"
            "selp.b64 bs, %7, %6, pd;\n\t"
            "mul.wide.u32 ad, t, 2;\n\t"
            "add.u64 ad, ad, bs;\n\t"
            "@po st.global.u16 [ad], val;\n\t"
            "}"
            : "+r"(list_ptr)
            : "r"(e), "r"(peek), "r"(qq), "r"(list_limit), "r"(slot0), "l"(coef_seg), "l"(dc_seg)
            : "memory");
#else
        put_symbol_generic(*this, e, peek, qq);
#endif
    }
};

// Policy of the repair path: tables in shared memory (LUT pool at s_dyn), the scan read from
// global memory.
struct GlobalMem {
    const SharedTables* st;
    const HuffSlow* slow_;
    uint32_t recs_addr;
    const uint32_t* words;
    uint32_t lut_addr;
    static constexpr uint32_t kWordStep = 1;
    __device__ __forceinline__ uint32_t lut(uint32_t tbl, uint32_t i) const { return lds32(tbl + i * 4u); }
    __device__ __forceinline__ uint32_t single_of(uint32_t tbl) const { return tbl + (uint32_t)kLut1Size * 4u; }
    // continuation field: 16-byte units from the start of the staged LUT pool (dynamic shared memory
    // behind the window; stage_tables<true> re-bases the field for its compact layout)
    __device__ __forceinline__ uint32_t cont_table(uint32_t e) const { return lut_addr + ((e >> 8) & 0x7fff0u); }
    template <bool WRITE>
    __device__ __forceinline__ uint32_t step(uint32_t e, uint32_t& pp, uint32_t& qq, uint32_t& ch,
                                             uint32_t& tbl, const SlotRec& r, uint32_t& w0, uint32_t& w1,
                                             uint32_t& w2, uint32_t& wn, uint32_t& peek, uint32_t end,
                                             uint32_t end_fast, uint32_t q_limit) const
    {
        return generic_step<WRITE>(*this, e, pp, qq, ch, tbl, r, w0, w1, w2, wn, peek, end, end_fast, q_limit);
    }
    __device__ __forceinline__ uint32_t slot_handle(uint32_t c) const
    {
        return recs_addr + c * (uint32_t)sizeof(SlotRec);
    }
    __device__ __forceinline__ SlotRec rec(uint32_t h) const { return lds_rec(h); }
    __device__ __forceinline__ uint32_t slot_tbls(uint32_t c) const { return st->tbls[c]; }
    __device__ __forceinline__ const HuffSlow& slow(uint32_t s) const { return slow_[s]; }
    __device__ __forceinline__ uint32_t word_handle(uint32_t i) const { return i; }
    __device__ __forceinline__ uint32_t word(uint32_t h) const { return __ldg(words + h); }
    __device__ __forceinline__ uint32_t probe(uint32_t tbl, uint32_t peek) const
    {
        return lut(tbl, peek >> (32 - kLutBits));
    }
    __device__ __forceinline__ uint32_t slide_peek(uint32_t& w0, uint32_t& w1, uint32_t& w2, uint32_t& wn,
                                                   uint32_t pp, uint32_t pn) const
    {
        slide_window_generic(*this, w0, w1, w2, wn, pp, pn);
        return funnel_left_wrap(w0, w1, pn);
    }
    __device__ __forceinline__ void note_long_code() const {}
    __device__ __forceinline__ void put_symbol(uint32_t, uint32_t, uint32_t) const {}
};

// Slice `block` of `nblocks` of a buffer, cleared by `nthreads` threads of which this is `tid`.
__device__ __forceinline__ void zero_fill(void* base, size_t bytes, uint32_t block, uint32_t nblocks,
                                          uint32_t tid, uint32_t nthreads)
{
    // bytes is a multiple of 16 (planner); every CTA clears one slice
    const size_t vecs = bytes / 16;
    const size_t per_cta = (vecs + nblocks - 1) / nblocks;
    const size_t begin = per_cta * block;
    const size_t end = begin + per_cta < vecs ? begin + per_cta : vecs;
    uint4* p = reinterpret_cast<uint4*>(base);
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (size_t i = begin + tid; i < end; i += nthreads) p[i] = z;
}

// The warp's share of huff_sync: fix-point iteration over its chunk, mailbox exchange with the
// chunk on its left, one record per owned subsequence.
__device__ __forceinline__ void sync_chunk(const ScanBufs& b, uint32_t num_subs, int num_segments,
                                           SharedTables& st, long long win_sub0, uint32_t chunk, int lane,
                                           int win_words)
{
    const long long sub = (long long)chunk * kChunkSubs - kOverlapSubs + lane;
    const uint32_t chunk_first = chunk * (uint32_t)kChunkSubs;
    const bool active = sub >= 0 && sub < (long long)num_subs;

    uint32_t k = 0, i = 0, seg_bits = 0;
    bool head = false;
    DevMem M{&st, b.slow, smem_addr(st.recs), 0, smem_addr(s_dyn + win_words / 4)};
    if (active) {
        k = find_segment(b.segs, num_segments, (uint32_t)sub);
        const SegDesc seg = b.segs[k];
        i = (uint32_t)sub - seg.sub_off;
        head = i == 0;
        seg_bits = seg.nbytes * 8u;
        M.win_addr = smem_addr(s_dyn) + (uint32_t)((int)((long long)seg.sub_off - win_sub0) * kSubBytes);
    }

    // fix-point iteration inside the warp; lanes <= frozen keep their entry state
    uint16_t entry = 0; // cold state == state at a segment start
    Mids mids;
#pragma unroll
    for (int k = 0; k < kWriteParts - 1; ++k) mids.v[k] = 0;
    unsigned long long rec = active ? decode_sub(M, seg_bits, i, entry, &mids) : 0ull;
    // One-entry memo: the evaluation this lane held before the current one.  A lane inside a
    // stretch that is slow to synchronise alternates between the state of the decoder that is
    // right and the state of one that is not yet (each arrives from the left one pass after the
    // other); switching back is then free, and the iteration needs as many decoding passes as an
    // oracle that picks the right neighbour state would (tools/sim/sync_sim.cpp: 18 -> 10 passes
    // for the slowest warp of the 12 MP workload, 34 -> 15 on IMG_6510.JPG).
    uint32_t memo_entry = 0xffffffffu; // no packed state has bit 16 set
    unsigned long long memo_rec = 0ull;
    Mids memo_mids = mids;
    auto adopt = [&](uint16_t want) {
        const uint32_t old_entry = entry;
        const unsigned long long old_rec = rec;
        const Mids old_mids = mids;
        entry = want;
        if (memo_entry == (uint32_t)want) {
            rec = memo_rec;
            mids = memo_mids;
        } else {
            rec = decode_sub(M, seg_bits, i, entry, &mids);
        }
        memo_entry = old_entry;
        memo_rec = old_rec;
        memo_mids = old_mids;
    };
    auto iterate = [&](int frozen) {
        for (;;) {
            const uint16_t px = (uint16_t)__shfl_up_sync(0xffffffffu, (uint32_t)entry_x(rec), 1);
            const uint16_t want = (lane <= frozen || head || !active) ? entry : px;
            const bool changed = want != entry;
            if (!__any_sync(0xffffffffu, changed)) break;
            if (changed) adopt(want);
        }
    };
    iterate(0);

    // mailbox: publish the exit state of the chunk's last subsequence, check the left neighbour.
    // A chunk reads its left mailbox once.  The value it read is final unless the left chunk
    // re-publishes a DIFFERENT exit state afterwards (its own entry changed and the change did not
    // die out inside its 26 subsequences), or the wait timed out: only then can a chunk boundary
    // disagree, and only then does the last CTA have to verify (and repair) the records.
    const uint32_t last_sub = umin(chunk_first + (uint32_t)kChunkSubs, num_subs) - 1u;
    const int last_lane = (int)(last_sub - chunk_first) + kOverlapSubs;
    volatile uint32_t* mailbox = b.chunk_state;
    const uint32_t first_exit = entry_x(rec);
    if (lane == last_lane) mailbox[chunk] = 0x80000000u | first_exit;
    if (chunk > 0) {
        uint32_t left = 0;
        if (lane == kOverlapSubs && !head) {
            for (int spin = 0; spin < kMailboxSpins; ++spin) {
                left = mailbox[chunk - 1];
                if (left) break;
                __nanosleep(100);
            }
            if (left == 0) atomicOr(&b.ctrl->need_verify, 1u); // went on with its own guess
        }
        left = __shfl_sync(0xffffffffu, left, kOverlapSubs);
        const uint16_t assumed = (uint16_t)__shfl_sync(0xffffffffu, (uint32_t)entry, kOverlapSubs);
        if (left != 0 && (uint16_t)(left & 0xffffu) != assumed) { // warp-uniform
            if (lane == kOverlapSubs) {
                adopt((uint16_t)(left & 0xffffu));
                atomicAdd(&b.ctrl->relinked, 1u);
            }
            iterate(kOverlapSubs);
            if (lane == last_lane && entry_x(rec) != first_exit) {
                mailbox[chunk] = 0x80000000u | entry_x(rec);
                atomicOr(&b.ctrl->need_verify, 1u); // the right neighbour may hold the old value
            }
        }
    }
    if (active && lane >= kOverlapSubs) {
        ulonglong2* out = reinterpret_cast<ulonglong2*>(b.entries) + (size_t)sub * (kRecWords / 2);
        out[0] = make_ulonglong2(
            pack_entry((uint16_t)(entry | (head ? kStateHead : 0)), entry_x(rec), entry_n(rec)),
            mids_word(mids, 0));
        if (kRecWords == 4) out[kRecWords / 2 - 1] = make_ulonglong2(mids_word(mids, kWriteParts == 4 ? 1 : 0), 0ull);
        b.seg_idx[sub] = k;
    }
}

// Runs in the CTA of huff_sync that finishes last: verify every chunk boundary (exit of the left
// record == entry of the right one).  If all agree, every record is exact by induction from the
// segment starts.  Otherwise repair sequentially from each break until the stored records agree
// again, and re-verify every boundary (concurrent repairs may interleave).  A mismatch is
// possible only when a mailbox wait timed out or a left neighbour re-published after it was
// read, so this is a few hundred loads in practice; on adversarial streams it degenerates to a
// serial decode, like the reference's TODO at src/decode_huffman.cu:475.
__device__ __forceinline__ void verify_and_repair(const ScanBufs& b, uint32_t N, SharedTables& st, int win_words)
{
    volatile unsigned long long* entries = reinterpret_cast<volatile unsigned long long*>(b.entries);
    uint32_t stride = (uint32_t)kChunkSubs; // first round: only chunk boundaries can disagree
    uint32_t rounds = 0, slow_total = 0;
    for (;;) {
        uint32_t bad = 0;
        for (uint32_t j = stride * (1u + threadIdx.x); j < N; j += stride * blockDim.x)
            bad |= boundary_ok(entries[kRecWords * (size_t)(j - 1)], entries[kRecWords * (size_t)j]) ? 0u : 1u;
        if (!__syncthreads_or((int)bad)) break;
        uint32_t rewritten = 0;
        for (uint32_t j = stride * (1u + threadIdx.x); j < N; j += stride * blockDim.x) {
            if (boundary_ok(entries[kRecWords * (size_t)(j - 1)], entries[kRecWords * (size_t)j])) continue;
            const SegDesc seg = b.segs[*(volatile uint32_t*)(b.seg_idx + j)];
            const GlobalMem M{&st, b.slow, smem_addr(st.recs),
                              b.destuffed + (size_t)seg.sub_off * kSubWords,
                              smem_addr(s_dyn + win_words / 4)};
            rewritten += relink_from(M, seg, entries, j);
        }
        slow_total += rewritten;
        ++rounds;
        stride = 1;
        __threadfence();
        __syncthreads();
    }
    if (slow_total) atomicAdd(&b.ctrl->slow_repairs, slow_total);
    if (threadIdx.x == 0) b.ctrl->link_rounds = rounds;
}

template <int WARPS>
__device__ __forceinline__ void huff_sync_body(const ScanGeom& g, const ScanBufs& b, const uint32_t block,
                                               const uint32_t nblocks)
{
    constexpr int kUseful = sync_useful(WARPS), kWinWords = sync_win_words(WARPS);
    __shared__ SharedTables st;
    __shared__ uint32_t s_last;
    __shared__ __align__(8) unsigned long long s_bar[2];
    const uint32_t bar_tables = JB2_BULK_STAGE ? smem_addr(&s_bar[0]) : 0u;
    const uint32_t bar_window = JB2_BULK_STAGE ? smem_addr(&s_bar[1]) : 0u;
    const unsigned long long t0 = now_ns();
    JB2_TL(b.ctrl, 2, true);
    const long long win_sub0 = (long long)block * kUseful - kOverlapSubs;
    // before the destuff kernel is done: tables into shared memory
    stage_tables<false>(g, b, st, kWinWords, bar_tables);
    if (JB2_FILL_WARP == 0 && !b.fill_in_write) {
        zero_fill(b.coef, b.coef_bytes, block, nblocks, threadIdx.x, blockDim.x);
        zero_fill(b.dc, b.dc_bytes, block, nblocks, threadIdx.x, blockDim.x);
    }
    if (JB2_FILL_WARP == 2 && (threadIdx.x >> 5) == WARPS && !b.fill_in_write) {
        zero_fill(b.coef, b.coef_bytes, block, nblocks, threadIdx.x & 31u, 32u);
        zero_fill(b.dc, b.dc_bytes, block, nblocks, threadIdx.x & 31u, 32u);
    }
    JB2_TL(b.ctrl, 20, false); // latest prologue done
    grid_dependency_wait();
    grid_dependency_trigger();
    JB2_TL(b.ctrl, 3, true);
    JB2_TL(b.ctrl, 4, false);
    // the totals of the scan come from the destuff kernel (the grid was sized from upper bounds)
    const uint32_t num_subs = __ldcg(&b.info->num_subs);
    const int num_segments = (int)__ldcg(&b.info->num_segments);
    stage_window(b, num_subs, win_sub0 * kSubWords, kWinWords, bar_window, bar_tables);
    const unsigned long long t1 = now_ns();
    JB2_TL(b.ctrl, 13, false);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t chunk = block * WARPS + warp;
    if (JB2_FILL_WARP == 2 && warp == WARPS) {
    } else if (JB2_FILL_WARP == 1 && warp == WARPS) {
        if (!b.fill_in_write) {
        // (mode 1: the extra warp fills while the other warps iterate)
        zero_fill(b.coef, b.coef_bytes, block, nblocks, (uint32_t)lane, 32u);
        zero_fill(b.dc, b.dc_bytes, block, nblocks, (uint32_t)lane, 32u);
        }
#ifdef JB2_TIMELINE
        if (lane == 0) tl_stamp(b.ctrl, 21, false);
#endif
    } else if (chunk * (uint32_t)kChunkSubs < num_subs) // warp-uniform
        sync_chunk(b, num_subs, num_segments, st, win_sub0, chunk, lane, kWinWords);
    const unsigned long long t2 = now_ns();

    // last CTA out verifies the records of the whole scan
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&b.ctrl->sync_done, 1u) == nblocks - 1u ? 1u : 0u;
    __syncthreads();
    JB2_TL(b.ctrl, 5, false);
    if (s_last) {
        __threadfence();
        if (kAlwaysVerify || __ldcg(&b.ctrl->need_verify)) verify_and_repair(b, num_subs, st, kWinWords); // block-uniform
        JB2_TL(b.ctrl, 14, false);
        // destuff is long done: re-arm its look-back descriptors and ticket for the next decode
        // that runs on this scratch without a new transfer (transfer uploads them as zeros)
        uint32_t* arm = reinterpret_cast<uint32_t*>(b.destuff_desc);
        for (uint32_t i = threadIdx.x; i < b.destuff_words; i += blockDim.x) arm[i] = 0u;
    }
    if (lane == 0) {
        atomicMax(&b.ctrl->sync_stage_ns, (uint32_t)(t1 - t0));
        atomicMax(&b.ctrl->sync_iter_ns, (uint32_t)(t2 - t1));
        if (s_last) b.ctrl->link_verify_ns = (uint32_t)(now_ns() - t2);
    }
    JB2_TL(b.ctrl, 6, false);
}

// The scan descriptors of a batched launch travel through shared memory (by-value kernel
// parameters live in the constant bank; a descriptor in global memory would be re-read at every use).
struct __align__(16) BatchScan {
    ScanGeom g;
    ScanBufs b;
};
__device__ __forceinline__ void load_batch_scan(BatchScan& dst, const BatchImage& im)
{
    static_assert(sizeof(ScanGeom) % 4 == 0 && sizeof(ScanBufs) % 4 == 0, "copied word by word");
    const uint32_t* sg = reinterpret_cast<const uint32_t*>(&im.g);
    const uint32_t* sb = reinterpret_cast<const uint32_t*>(&im.b);
    uint32_t* dg = reinterpret_cast<uint32_t*>(&dst.g);
    uint32_t* db = reinterpret_cast<uint32_t*>(&dst.b);
    for (uint32_t i = threadIdx.x; i < sizeof(ScanGeom) / 4; i += blockDim.x) dg[i] = __ldg(sg + i);
    for (uint32_t i = threadIdx.x; i < sizeof(ScanBufs) / 4; i += blockDim.x) db[i] = __ldg(sb + i);
    __syncthreads();
}

__global__ void __launch_bounds__(kSyncThreads, 1)
huff_sync(const ScanGeom g, const ScanBufs b)
{
    huff_sync_body<kSyncWarps>(g, b, blockIdx.x, gridDim.x);
}

__global__ void __launch_bounds__(kSyncThreadsBatch, 1)
huff_sync_batch(const BatchImage* __restrict__ images)
{
    const BatchImage& im = images[blockIdx.y];
    const uint32_t nblocks = im.ctas[kBkSync];
    if (blockIdx.x >= nblocks) return;
    __shared__ BatchScan sc;
    load_batch_scan(sc, im);
    huff_sync_body<kSyncWarpsBatch>(sc.g, sc.b, blockIdx.x, nblocks);
}

__device__ __forceinline__ unsigned long long ld_acquire64(const unsigned long long* p)
{
    unsigned long long v;
    // relaxed: the word itself carries the payload, nothing else is read behind it
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release64(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// First output slot (segment relative) of the CTA's subsequences == segmented exclusive prefix sum
// of the slot counts of all records before them (replaces the reference's CUB ExclusiveScanByKey +
// assign_sinfo_n, src/decode_huffman.cu:827-869, and a whole kernel of ours).  Every CTA scans
// its own records, publishes {slots after its last segment head, has-head} in one 64-bit word and
// adds up the words of the CTAs before it, back to the nearest one that holds a segment head.
// The published words do not depend on other CTAs, so nothing chains; CTAs are numbered by an
// atomic ticket, so a CTA only ever waits for CTAs that are already running.
// value: slot count (even threads: the record's n, odd threads: 0); returns the exclusive prefix.
__device__ __forceinline__ uint32_t cta_slot_prefix(uint32_t cta, uint32_t n, bool head,
                                                    unsigned long long* agg)
{
    __shared__ uint32_t s_sum[kWriteThreads / 32], s_flag[kWriteThreads / 32];
    __shared__ uint32_t s_excl[kWriteThreads / 32], s_seen[kWriteThreads / 32];
    __shared__ uint32_t s_min[kWriteThreads / 32], s_add[kWriteThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    // inclusive segmented scan over the threads of the CTA
    uint32_t inc = n, f = head ? 1u : 0u; // sum of n from the nearest head (inclusive) to here
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, inc, d);
        const uint32_t upf = __shfl_up_sync(0xffffffffu, f, d);
        if (lane >= d) {
            if (!f) inc += up;
            f |= upf;
        }
    }
    if (lane == 31) s_sum[warp] = inc, s_flag[warp] = f;
    __syncthreads();
    if (warp == 0) {
        uint32_t v = lane < warps ? s_sum[lane] : 0u, gf = lane < warps ? s_flag[lane] : 0u;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, v, d);
            const uint32_t upg = __shfl_up_sync(0xffffffffu, gf, d);
            if (lane >= d) {
                if (!gf) v += up;
                gf |= upg;
            }
        }
        // exclusive over warps: value and "a head was seen" before warp `lane`
        const uint32_t ev = __shfl_up_sync(0xffffffffu, v, 1), eg = __shfl_up_sync(0xffffffffu, gf, 1);
        if (lane < warps) {
            s_excl[lane] = lane == 0 ? 0u : ev;
            s_seen[lane] = lane == 0 ? 0u : eg;
        }
        if (lane == warps - 1) // CTA aggregate: slots after the last head, has-head, ready
            st_release64(agg + cta, ((unsigned long long)v << 2) | (gf ? 2ull : 0ull) | 1ull);
    }
    __syncthreads();
    // inclusive value / head-seen flag up to and including this thread, inside the CTA
    const uint32_t local_inc = f ? inc : inc + s_excl[warp];
    const bool local_seen = f != 0 || s_seen[warp] != 0;

    // look back over the CTAs before this one
    uint32_t carry = 0;
    for (uint32_t base = 0; base < cta; base += blockDim.x) {
        const uint32_t t = base + threadIdx.x;
        unsigned long long w = 0;
        if (t < cta) {
            const unsigned long long* src = agg + (cta - 1u - t);
            while (((w = ld_acquire64(src)) & 1ull) == 0) __nanosleep(40);
        }
        const uint32_t mine = (w & 2ull) ? t : 0xffffffffu;
        const uint32_t wmin = __reduce_min_sync(0xffffffffu, mine);
        if (lane == 0) s_min[warp] = wmin;
        __syncthreads();
        uint32_t tmin = 0xffffffffu;
        for (int k = 0; k < warps; ++k) tmin = umin(tmin, s_min[k]);
        const uint32_t part = __reduce_add_sync(0xffffffffu, (t < cta && t <= tmin) ? (uint32_t)(w >> 2) : 0u);
        if (lane == 0) s_add[warp] = part;
        __syncthreads();
        for (int k = 0; k < warps; ++k) carry += s_add[k];
        if (tmin != 0xffffffffu) break; // uniform
    }
    const uint32_t incl = local_seen ? local_inc : local_inc + carry;
    return head ? 0u : incl - n;
}

// Zero coefficient slots [A, B) (int16 indices into the coefficient buffer) and the DC entries of
// the data units that begin inside, with the whole warp (A and B are warp-uniform).
__device__ __forceinline__ void warp_zero_slots(int16_t* __restrict__ coef, int16_t* __restrict__ dc,
                                                uint32_t A, uint32_t B, uint32_t lane)
{
    if (A >= B) return;
    for (uint32_t du = (A + 63u) / 64u + lane; du < (B + 63u) / 64u; du += 32u) dc[du] = 0;
    const uint32_t A8 = (A + 7u) & ~7u, B8 = B & ~7u; // whole 16-byte vectors: [A8, B8)
    if (A8 >= B8) {
        for (uint32_t s = A + lane; s < B; s += 32u) coef[s] = 0;
        return;
    }
    if (A + lane < A8) coef[A + lane] = 0;
    if (B8 + lane < B) coef[B8 + lane] = 0;
    uint4* p = reinterpret_cast<uint4*>(coef);
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (uint32_t v = (A8 >> 3) + lane; v < (B8 >> 3); v += 32u) p[v] = z;
}

__device__ __forceinline__ void huff_write_body(const ScanGeom& g, const ScanBufs& b)
{
    __shared__ SharedTables st;
    __shared__ uint32_t s_cta;
    __shared__ __align__(8) unsigned long long s_bar;
    const unsigned long long t0 = now_ns();
    JB2_TL(b.ctrl, 7, true);
    const uint32_t subs_per_cta = blockDim.x / kWriteParts;
    const int win_words = (int)(subs_per_cta * kSubWords + kWinMargin);
    // before huff_sync is done (the control block was cleared two kernels ago): ticket, tables
    if (threadIdx.x == 0) s_cta = atomicAdd(&b.ctrl->write_ticket, 1u);
    stage_tables<true>(g, b, st, win_words);
    __syncthreads();
    // The scan itself comes from the destuff kernel, the grand-predecessor: it is complete when this
    // CTA starts (huff_sync triggers its dependents only behind its own dependency wait), so the
    // window is staged while huff_sync is still at work; only the records have to wait.
    const uint32_t num_subs = __ldcg(&b.info->num_subs);
    const uint32_t cta = s_cta;
    const uint32_t sub0 = cta * subs_per_cta;
    stage_window(b, num_subs, (long long)sub0 * kSubWords, win_words, JB2_BULK_STAGE ? smem_addr(&s_bar) : 0u);
    grid_dependency_wait();
    grid_dependency_trigger();
    JB2_TL(b.ctrl, 15, true);
    JB2_TL(b.ctrl, 8, false);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t j = sub0 + threadIdx.x / kWriteParts, h = threadIdx.x % kWriteParts;

    // this thread's record; the CTA's slot prefix is published before the tables are staged, so
    // that the CTAs after this one never wait for it
    unsigned long long rec = 0;
    uint32_t mid = 0, mid_next = 0, k = 0; // state at this part's start / at its end (h < last)
    if (j < num_subs) {
        const ulonglong2* rp = reinterpret_cast<const ulonglong2*>(b.entries) + (size_t)j * (kRecWords / 2);
        const ulonglong2 r = __ldcg(rp);
        rec = r.x;
        unsigned long long w[kRecWords - 1];
        w[0] = r.y;
        if (kRecWords == 4) {
            const ulonglong2 r2 = __ldcg(rp + 1);
            w[1] = r2.x;
            w[kRecWords - 2] = r2.y;
        }
        // boundary states by select (a dynamically indexed array would live in local memory)
#pragma unroll
        for (int kk = 0; kk < kWriteParts - 1; ++kk) {
            const uint32_t m = mid_of(w, kk);
            mid = h == (uint32_t)kk + 1u ? m : mid;
            mid_next = h == (uint32_t)kk ? m : mid_next;
        }
        k = __ldcg(b.seg_idx + j);
    }
    const bool is_head = j < num_subs && (entry_g(rec) & kStateHead) != 0;
    uint32_t q0 = cta_slot_prefix(cta, h == 0 ? entry_n(rec) : 0u, h == 0 && is_head, b.write_agg);
    q0 = __shfl_sync(0xffffffffu, q0, lane & ~(kWriteParts - 1)); // the sub's first thread has it
    if (h == 0 && j < num_subs) b.qpos[j] = q0; // kept for inspection (jpeggpu_b200_scan_stats)

    const unsigned long long t1 = now_ns();

    // this thread's half subsequence: placement, slot range [qs, qe)
    bool has = false;
    uint32_t first_du = 0, q_limit = 0, qs = 0, qe = 0, i = 0;
    SegDesc seg = {0, 0, 0, 0};
    if (j < num_subs) {
        seg = b.segs[k];
        i = j - seg.sub_off;
        if (segment_extent(g, k, first_du, q_limit)) {
            qs = h ? q0 + (mid >> 16) : q0;
            qe = umin(h + 1 < (uint32_t)kWriteParts ? q0 + (mid_next >> 16) : q0 + entry_n(rec), q_limit);
            has = qe > qs;
        }
    }
    // Zero-fill of the output by the writers themselves, for buffers that do not fit L2.  The slot ranges of consecutive threads
    // tile the coefficient buffer (the slot prefix sum says where each one starts), so every warp
    // clears exactly the stretch its 32 threads are about to scatter into - 20 KB of coalesced
    // stores - and the lines are still in L2 when the scattered 2-byte stores arrive.  Clearing the
    // whole buffer up front (huff_sync's extra warp) is cheaper while the buffer fits L2 (12 MP:
    // decode 128.7 us against 132.1 with this path), but at 48 MP (146 MB) every scattered store
    // became a DRAM read-modify-write: 354 -> 320 us with this path.
    // What no thread will write is cleared too: the last thread of a restart segment clears up to
    // the first data unit of the next segment that holds data (or the end of the buffer), the very
    // first thread from the start of the buffer.
    if (b.fill_in_write) { // block-uniform
        uint32_t fs = 0, fe = 0; // this thread's stretch, in global slots
        bool mine = false;
        if (j < num_subs && q_limit != 0) {
            const uint32_t base = first_du * 64u;
            const uint32_t q_begin = h ? q0 + (mid >> 16) : q0;
            const uint32_t q_end = h + 1 < (uint32_t)kWriteParts ? q0 + (mid_next >> 16) : q0 + entry_n(rec);
            fs = base + umin(q_begin, q_limit);
            fe = base + umin(q_end, q_limit);
            if (j == 0 && h == 0) fs = 0;
            const uint32_t seg_subs = (seg.nbytes + (uint32_t)kSubBytes - 1u) / (uint32_t)kSubBytes;
            if (i + 1 == seg_subs && h + 1 == (uint32_t)kWriteParts) {
                // last thread of the segment: up to the next segment that holds data
                const uint32_t nseg = __ldcg(&b.info->num_segments);
                uint32_t k2 = k + 1, du2 = 0, lim2 = 0;
                while (k2 < nseg && b.segs[k2].nbytes == 0) ++k2;
                fe = (k2 < nseg && segment_extent(g, k2, du2, lim2)) ? du2 * 64u : (uint32_t)(b.coef_bytes / 2);
            }
            mine = true;
        }
        const uint32_t next_fs = __shfl_down_sync(0xffffffffu, fs, 1);
        const bool chained = mine && (lane == 31 || fe == next_fs);
        if (__all_sync(0xffffffffu, chained)) {
            warp_zero_slots(b.coef, b.dc, __shfl_sync(0xffffffffu, fs, 0), __shfl_sync(0xffffffffu, fe, 31),
                            (uint32_t)lane);
        } else {
            // a warp with idle threads (end of the scan) or a surplus segment: stretch by stretch
            for (int l = 0; l < 32; ++l)
                warp_zero_slots(b.coef, b.dc, __shfl_sync(0xffffffffu, fs, l), __shfl_sync(0xffffffffu, fe, l),
                                (uint32_t)lane);
        }
        __syncwarp();
    }
    const uint32_t lut_vec = (g.lut_compact_words + 3) / 4;
    uint8_t* warp_mem = reinterpret_cast<uint8_t*>(s_dyn + win_words / 4 + lut_vec) + (size_t)warp * kWarpSmem;
    const uint32_t list_addr = smem_addr(warp_mem) + (uint32_t)lane * (kListStride * 4u);
    uint32_t list_end = list_addr;
    if (has) {
        WriteMem M;
        M.st = &st;
        M.slow_ = b.slow;
        M.recs_addr = smem_addr(st.recs);
        M.lut_addr = smem_addr(s_dyn + win_words / 4);
        M.win_addr = smem_addr(s_dyn) + (uint32_t)((int)((long long)seg.sub_off - (long long)sub0) * kSubBytes);
        M.list_ptr = list_addr;
        M.list_limit = list_addr + (uint32_t)kListCap * 4u;
        M.slot0 = qs;
        M.coef_seg = b.coef + (size_t)first_du * 64;
        M.dc_seg = b.dc + first_du;
        write_part(M, seg, i, h, rec, mid, q0, q_limit);
        list_end = M.list_ptr;
    }
    const unsigned long long t2 = now_ns();

    // Every thread scatters its own list straight into the buffers huff_sync zero-filled: AC
    // values at their zig-zag slot, DC differences into the compact per-DU array.  (Assembling
    // whole data units in shared memory and storing 16-byte vectors was 15% slower: ~3000 more
    // warp instructions per warp and, through its shared-memory tiles, one wave more of CTAs.)
    const uint32_t gs0 = first_du * 64u + qs; // global slot of this thread's first coefficient
#ifndef JB2_WRITE_MERGE
#define JB2_WRITE_MERGE 1
#endif
    // Two list entries that fill one aligned 32-bit word of the coefficient buffer (an even
    // zig-zag slot and its successor - the low frequencies are usually all non-zero) leave as one
    // store: scattered stores cost one L1 sector slot each, whatever their width.  Only words
    // whose two halves both belong to THIS thread's list are merged, so no other thread's
    // coefficient can be overwritten by a zero half.
    for (uint32_t at = list_addr; at != list_end;) {
        const uint32_t e = lds32(at);
        const uint32_t slot = gs0 + (e >> 16);
        at += 4u;
        if ((slot & 63u) == 0) {
            b.dc[slot >> 6] = (int16_t)(e & 0xffffu);
            continue;
        }
        if (JB2_WRITE_MERGE && (slot & 1u) == 0 && at != list_end) {
            const uint32_t e2 = lds32(at);
            if (gs0 + (e2 >> 16) == slot + 1u) {
                *reinterpret_cast<uint32_t*>(b.coef + slot) = (e & 0xffffu) | (e2 << 16);
                at += 4u;
                continue;
            }
        }
        b.coef[slot] = (int16_t)(e & 0xffffu);
    }
    if (lane == 0) {
        atomicMax(&b.ctrl->write_stage_ns, (uint32_t)(t1 - t0));
        atomicMax(&b.ctrl->write_decode_ns, (uint32_t)(t2 - t1));
        atomicMax(&b.ctrl->write_flush_ns, (uint32_t)(now_ns() - t2));
    }
    JB2_TL(b.ctrl, 9, false);
}

__global__ void __launch_bounds__(kWriteThreads, 2)
huff_write(const ScanGeom g, const ScanBufs b)
{
    huff_write_body(g, b);
}

__global__ void __launch_bounds__(kWriteThreads, 2)
huff_write_batch(const BatchImage* __restrict__ images)
{
    const BatchImage& im = images[blockIdx.y];
    if (blockIdx.x >= im.ctas[kBkWrite]) return;
    __shared__ BatchScan sc;
    load_batch_scan(sc, im);
    huff_write_body(sc.g, sc.b);
}

} // namespace

namespace {
size_t write_smem_bytes(int warps, size_t lut_write_bytes)
{
    return (size_t)(warps * kSubsPerWarp * kSubWords + kWinMargin) * 4 + lut_write_bytes + (size_t)warps * kWarpSmem;
}
void configure_huffman()
{
    static PerDeviceOnce configured;
    configured.run([] {
        const int max_smem = 226 * 1024; // 227 KiB per CTA minus the static part
        prefer_max_shared(huff_sync, max_smem);
        prefer_max_shared(huff_write, max_smem);
        prefer_max_shared(huff_sync_batch, max_smem);
        prefer_max_shared(huff_write_batch, max_smem);
    });
}
} // namespace

// Warps per huff_write CTA: as many as let TWO CTAs share an SM (each warp needs its coefficient
// lists, each CTA its own copy of the tables); with tables too large for that, as many as fit one.
int batch_write_warps(uint32_t lut_compact_words_max)
{
    const size_t lut_write_bytes = (size_t)((lut_compact_words_max + 3) / 4) * 16;
    const size_t per_sm = 228 * 1024, reserved = 1024 + 1024; // per-CTA reservation + static shared memory
    auto fit = [&](size_t budget) {
        int warps = kWriteThreads / 32;
        while (warps > 0 && write_smem_bytes(warps, lut_write_bytes) + reserved > budget) --warps;
        return warps;
    };
    const int two = fit(per_sm / 2);
    if (two >= 4) return two;
    const int one = fit(227 * 1024);
    return one > 0 ? one : 1;
}

// One image alone: a small image gets smaller CTAs so that its write threads spread over all SMs
// (2 MP: 25 CTAs of 10 warps left 123 SMs idle; huff_write 32.8 -> 26.6 us at 2 MP, 33.0 -> 30.9 at
// 6 MP, unchanged from 12 MP on; a minimum of 2 warps instead of 4 measured 28.9 / 32.6).
static int single_write_warps(uint32_t num_subs, uint32_t lut_compact_words)
{
    const int cap = batch_write_warps(lut_compact_words);
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const uint32_t total_warps = (num_subs * (uint32_t)kWriteParts + 31u) / 32u;
    const uint32_t slots = 2u * (uint32_t)(sms > 0 ? sms : 148); // two CTAs per SM
    int want = (int)((total_warps + slots - 1u) / slots);
    if (want < JB2_WRITE_MIN_WARPS) want = JB2_WRITE_MIN_WARPS;
    return want < cap ? want : cap;
}

uint32_t huff_sync_ctas(uint32_t num_subs, bool batch)
{
    const uint32_t chunks = (num_subs + kChunkSubs - 1) / kChunkSubs;
    const uint32_t warps = batch ? kSyncWarpsBatch : kSyncWarps;
    return (chunks + warps - 1) / warps;
}
uint32_t huff_write_ctas(uint32_t num_subs, int warps)
{
    const uint32_t subs_per_cta = (uint32_t)warps * (uint32_t)kSubsPerWarp;
    return (num_subs + subs_per_cta - 1) / subs_per_cta;
}

cudaError_t launch_huffman_batch(const BatchLaunch& bl, int write_warps, cudaStream_t stream)
{
    configure_huffman();
    const size_t lut_bytes = (size_t)((bl.lut_pool_words_max + 3) / 4) * 16;
    const size_t lut_write_bytes = (size_t)((bl.lut_compact_words_max + 3) / 4) * 16;
    JB2_LAUNCH(launch_dependent(huff_sync_batch, dim3(bl.max_ctas[kBkSync], bl.count), kSyncThreadsBatch,
                                sync_win_words(kSyncWarpsBatch) * 4 + lut_bytes, stream, bl.d_images));
    return launch_dependent(huff_write_batch, dim3(bl.max_ctas[kBkWrite], bl.count), write_warps * 32,
                            write_smem_bytes(write_warps, lut_write_bytes), stream, bl.d_images);
}

cudaError_t launch_huffman(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm)
{
    if (g.num_subs == 0) return cudaSuccess;
    configure_huffman();
    const size_t lut_bytes = (size_t)((g.lut_pool_words + 3) / 4) * 16;
    const size_t lut_write_bytes = (size_t)((g.lut_compact_words + 3) / 4) * 16;
    tm->mark(stream, kIdSync);
    JB2_LAUNCH(launch_dependent(huff_sync, huff_sync_ctas(g.num_subs, false), kSyncThreads,
                                sync_win_words(kSyncWarps) * 4 + lut_bytes, stream, g, b));
    const int warps = single_write_warps(g.num_subs, g.lut_compact_words);
    tm->mark(stream, kIdWrite);
    return launch_dependent(huff_write, huff_write_ctas(g.num_subs, warps), warps * 32,
                            write_smem_bytes(warps, lut_write_bytes), stream, g, b);
}

} // namespace jb2
