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
//               corrected state (rare).  One 8-byte record {entry, exit, slots} per subsequence.
//               The kernel also zero-fills the coefficient buffer (its load/store units are idle
//               while the warps decode).
//   huff_link   one CTA: verifies every chunk boundary (exit of the left record == entry of the
//               right one).  If all agree, every record is exact by induction from the segment
//               starts.  Otherwise it repairs sequentially (degenerates to a serial decode on
//               adversarial streams, like the reference's TODO at src/decode_huffman.cu:475).
//               Then a segmented exclusive scan of the slot counts gives the first output slot of
//               every subsequence (replaces the CUB ExclusiveScanByKey + assign_sinfo_n,
//               src/decode_huffman.cu:827-869).
//   huff_write  decode from the verified entry states; scatter non-zero AC coefficients into the
//               zeroed DU-major buffer and DC differences into a compact per-DU array
//               (src/decode_huffman.cu:627-682).
#include "huff_threads.hpp"
#include "kernels.hpp"

namespace jb2 {
namespace {

__constant__ uint8_t c_natural[64] = {JB2_NATURAL_ORDER};

constexpr int kSyncWarps = kHuffThreads / 32;
constexpr int kSyncUseful = kSyncWarps * kChunkSubs;            // records written per CTA
constexpr int kSyncWinSubs = kSyncUseful + kOverlapSubs;        // subsequences staged per CTA
constexpr int kWinMargin = 4;                                   // words past the last subsequence
constexpr int kSyncWinWords = kSyncWinSubs * kSubWords + kWinMargin;
constexpr int kWriteWinWords = kHuffThreads * kSubWords + kWinMargin;
static_assert(kSyncWinWords % 4 == 0 && kWriteWinWords % 4 == 0, "window must be uint4 sized");

struct SharedTables {
    SlotPair pairs[kMaxDuPerMcu];
    uint32_t tbls[kMaxDuPerMcu];
    uint8_t natural[64];
};

extern __shared__ __align__(16) uint4 s_dyn[]; // [window words | LUT pool]

// Stage the LUT pool, the slot records and words [first_word, first_word + win_words) of the
// destuffed scan (zero outside [0, total_words)).
template <int WIN_WORDS>
__device__ __forceinline__ void stage(const ScanGeom& g, const ScanBufs& b, SharedTables& st,
                                      long long first_word)
{
    uint32_t* win = reinterpret_cast<uint32_t*>(s_dyn);
    const long long total_words = (long long)g.num_subs * kSubWords;
    const uint4* src = reinterpret_cast<const uint4*>(b.destuffed);
    for (int i = threadIdx.x; i < WIN_WORDS / 4; i += blockDim.x) {
        const long long w = first_word + 4ll * i;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (w >= 0 && w + 4 <= total_words) v = __ldg(src + (w >> 2));
        reinterpret_cast<uint4*>(win)[i] = v;
    }
    const uint32_t lut_vec = (g.lut_pool_u16 + 7) / 8; // pool is padded to 16 B by the planner
    const uint4* lsrc = reinterpret_cast<const uint4*>(b.lut_pool);
    uint4* ldst = s_dyn + WIN_WORDS / 4;
    for (uint32_t i = threadIdx.x; i < lut_vec; i += blockDim.x) ldst[i] = __ldg(lsrc + i);
    if (threadIdx.x == 0) make_slot_pairs(g, st.pairs, st.tbls);
    if (threadIdx.x >= 32 && threadIdx.x < 96) st.natural[threadIdx.x - 32] = c_natural[threadIdx.x - 32];
    __syncthreads();
}

// Device memory policy of decode_span over the staged window.  `wrel` maps a segment-relative
// word index to a window index.
template <int WIN_WORDS>
struct DevMem {
    const SharedTables* st;
    const HuffSlow* slow_;
    int wrel;
    __device__ __forceinline__ uint16_t lut(uint32_t i) const
    {
        return reinterpret_cast<const uint16_t*>(s_dyn + WIN_WORDS / 4)[i];
    }
    __device__ __forceinline__ SlotPair pair(uint32_t c) const
    {
        const uint2 v = *reinterpret_cast<const uint2*>(&st->pairs[c]);
        return SlotPair{v.x, v.y};
    }
    __device__ __forceinline__ uint32_t slot_tbls(uint32_t c) const { return st->tbls[c]; }
    __device__ __forceinline__ const HuffSlow& slow(uint32_t s) const { return slow_[s]; }
    __device__ __forceinline__ uint32_t natural(uint32_t i) const { return st->natural[i]; }
    __device__ __forceinline__ uint32_t word(uint32_t i) const
    {
        return reinterpret_cast<const uint32_t*>(s_dyn)[(int)i + wrel];
    }
};

// Fallback policy of huff_link: tables in shared memory, the scan read from global memory.
struct GlobalMem {
    const SharedTables* st;
    const HuffSlow* slow_;
    const uint32_t* words;
    __device__ __forceinline__ uint16_t lut(uint32_t i) const
    {
        return reinterpret_cast<const uint16_t*>(s_dyn)[i];
    }
    __device__ __forceinline__ SlotPair pair(uint32_t c) const { return st->pairs[c]; }
    __device__ __forceinline__ uint32_t slot_tbls(uint32_t c) const { return st->tbls[c]; }
    __device__ __forceinline__ const HuffSlow& slow(uint32_t s) const { return slow_[s]; }
    __device__ __forceinline__ uint32_t natural(uint32_t i) const { return st->natural[i]; }
    __device__ __forceinline__ uint32_t word(uint32_t i) const { return __ldg(words + i); }
};

__device__ __forceinline__ void zero_fill(void* base, size_t bytes)
{
    // bytes is a multiple of 16 (planner); every CTA clears one slice
    const size_t vecs = bytes / 16;
    const size_t per_cta = (vecs + gridDim.x - 1) / gridDim.x;
    const size_t begin = per_cta * blockIdx.x;
    const size_t end = begin + per_cta < vecs ? begin + per_cta : vecs;
    uint4* p = reinterpret_cast<uint4*>(base);
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (size_t i = begin + threadIdx.x; i < end; i += blockDim.x) p[i] = z;
}

__global__ void __launch_bounds__(kHuffThreads)
huff_sync(const ScanGeom g, const ScanBufs b)
{
    __shared__ SharedTables st;
    const long long win_sub0 = (long long)blockIdx.x * kSyncUseful - kOverlapSubs;
    stage<kSyncWinWords>(g, b, st, win_sub0 * kSubWords);
    zero_fill(b.coef, b.coef_bytes);
    zero_fill(b.dc, b.dc_bytes);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t chunk = blockIdx.x * kSyncWarps + warp;
    const long long sub = (long long)chunk * kChunkSubs - kOverlapSubs + lane;
    const uint32_t chunk_first = chunk * (uint32_t)kChunkSubs;
    if (chunk_first >= g.num_subs) return; // whole warp
    const bool active = sub >= 0 && sub < (long long)g.num_subs;

    uint32_t k = 0, i = 0, seg_bits = 0;
    bool head = false;
    DevMem<kSyncWinWords> M{&st, b.slow, 0};
    if (active) {
        k = find_segment(b.segs, g.num_segments, (uint32_t)sub);
        const SegDesc seg = b.segs[k];
        i = (uint32_t)sub - seg.sub_off;
        head = i == 0;
        seg_bits = seg.nbytes * 8u;
        M.wrel = (int)((long long)seg.sub_off - win_sub0) * kSubWords;
    }

    // fix-point iteration inside the warp; lanes <= frozen keep their entry state
    uint16_t entry = 0; // cold state == state at a segment start
    unsigned long long rec = active ? decode_sub(M, seg_bits, i, entry) : 0ull;
    auto iterate = [&](int frozen) {
        for (;;) {
            const uint16_t px = (uint16_t)__shfl_up_sync(0xffffffffu, (uint32_t)entry_x(rec), 1);
            const uint16_t want = (lane <= frozen || head || !active) ? entry : px;
            const bool changed = want != entry;
            if (!__any_sync(0xffffffffu, changed)) break;
            if (changed) {
                entry = want;
                rec = decode_sub(M, seg_bits, i, entry);
            }
        }
    };
    iterate(0);

    // mailbox: publish the exit state of the chunk's last subsequence, check the left neighbour
    const uint32_t last_sub = umin(chunk_first + (uint32_t)kChunkSubs, g.num_subs) - 1u;
    const int last_lane = (int)(last_sub - chunk_first) + kOverlapSubs;
    volatile uint32_t* mailbox = b.chunk_state;
    if (lane == last_lane) mailbox[chunk] = 0x80000000u | entry_x(rec);
    if (chunk > 0) {
        uint32_t left = 0;
        if (lane == kOverlapSubs && !head) {
            for (int spin = 0; spin < (1 << 18); ++spin) {
                left = mailbox[chunk - 1];
                if (left) break;
                __nanosleep(100);
            }
        }
        left = __shfl_sync(0xffffffffu, left, kOverlapSubs);
        const uint16_t assumed = (uint16_t)__shfl_sync(0xffffffffu, (uint32_t)entry, kOverlapSubs);
        if (left != 0 && (uint16_t)(left & 0xffffu) != assumed) { // warp-uniform
            if (lane == kOverlapSubs) {
                entry = (uint16_t)(left & 0xffffu);
                rec = decode_sub(M, seg_bits, i, entry);
                atomicAdd(&b.ctrl->relinked, 1u);
            }
            iterate(kOverlapSubs);
            if (lane == last_lane) mailbox[chunk] = 0x80000000u | entry_x(rec);
        }
    }
    if (active && lane >= kOverlapSubs) {
        reinterpret_cast<unsigned long long*>(b.entries)[sub] =
            pack_entry((uint16_t)(entry | (head ? kStateHead : 0)), entry_x(rec), entry_n(rec));
        b.seg_idx[sub] = k;
    }
}

constexpr int kLinkThreads = 1024;

__global__ void __launch_bounds__(kLinkThreads)
huff_link(const ScanGeom g, const ScanBufs b)
{
    __shared__ SharedTables st;
    __shared__ uint32_t s_sum[kLinkThreads / 32], s_flag[kLinkThreads / 32];
    volatile unsigned long long* entries = reinterpret_cast<volatile unsigned long long*>(b.entries);
    const uint32_t N = g.num_subs;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    // ---- verify (and, rarely, repair) ------------------------------------------------------
    uint32_t stride = (uint32_t)kChunkSubs; // first round: only chunk boundaries can disagree
    uint32_t rounds = 0, slow_total = 0;
    bool staged = false;
    for (;;) {
        uint32_t bad = 0;
        for (uint32_t j = stride * (1u + threadIdx.x); j < N; j += stride * kLinkThreads)
            bad |= boundary_ok(entries[j - 1], entries[j]) ? 0u : 1u;
        if (!__syncthreads_or((int)bad)) break;
        if (!staged) { // tables are only needed on this path
            const uint32_t lut_vec = (g.lut_pool_u16 + 7) / 8;
            for (uint32_t v = threadIdx.x; v < lut_vec; v += kLinkThreads)
                s_dyn[v] = __ldg(reinterpret_cast<const uint4*>(b.lut_pool) + v);
            if (threadIdx.x == 0) make_slot_pairs(g, st.pairs, st.tbls);
            if (threadIdx.x < 64) st.natural[threadIdx.x] = c_natural[threadIdx.x];
            staged = true;
            __syncthreads();
        }
        uint32_t rewritten = 0;
        for (uint32_t j = stride * (1u + threadIdx.x); j < N; j += stride * kLinkThreads) {
            if (boundary_ok(entries[j - 1], entries[j])) continue;
            const SegDesc seg = b.segs[b.seg_idx[j]];
            const GlobalMem M{&st, b.slow, b.destuffed + (size_t)seg.sub_off * kSubWords};
            rewritten += relink_from(M, seg, entries, j);
        }
        slow_total += rewritten;
        ++rounds;
        stride = 1; // concurrent repairs may interleave: from now on check every boundary
        __syncthreads();
    }

    // ---- segmented exclusive scan of the slot counts ---------------------------------------
    const uint32_t per = (N + kLinkThreads - 1) / kLinkThreads;
    const uint32_t j0 = umin(per * threadIdx.x, N), j1 = umin(j0 + per, N);
    uint32_t sum = 0, flag = 0;
    for (uint32_t j = j0; j < j1; ++j) {
        const unsigned long long e = entries[j];
        if (entry_g(e) & kStateHead) sum = 0, flag = 1;
        sum += entry_n(e);
    }
    // inclusive segmented scan over threads of (sum, flag)
    uint32_t inc = sum, f = flag;
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
    uint32_t carry = 0;
    if (!f)
        for (int w = warp - 1; w >= 0; --w) {
            carry += s_sum[w];
            if (s_flag[w]) break;
        }
    inc += carry; // inclusive over threads, restarted at segment heads
    // Carry-in of thread t = inclusive value of thread t-1 (segment restarts already applied).
    __shared__ uint32_t s_inc[kLinkThreads];
    s_inc[threadIdx.x] = inc;
    __syncthreads();
    uint32_t acc = threadIdx.x ? s_inc[threadIdx.x - 1] : 0u;
    for (uint32_t j = j0; j < j1; ++j) {
        const unsigned long long e = entries[j];
        if (entry_g(e) & kStateHead) acc = 0;
        b.qpos[j] = acc;
        acc += entry_n(e);
    }
    if (slow_total) atomicAdd(&b.ctrl->slow_repairs, slow_total);
    if (threadIdx.x == 0) b.ctrl->link_rounds = rounds;
}

__global__ void __launch_bounds__(kHuffThreads)
huff_write(const ScanGeom g, const ScanBufs b)
{
    __shared__ SharedTables st;
    const uint32_t sub0 = blockIdx.x * kHuffThreads;
    stage<kWriteWinWords>(g, b, st, (long long)sub0 * kSubWords);
    const uint32_t j = sub0 + threadIdx.x;
    if (j >= g.num_subs) return;
    const uint32_t k = b.seg_idx[j];
    const SegDesc seg = b.segs[k];
    const DevMem<kWriteWinWords> M{&st, b.slow, (int)((long long)seg.sub_off - (long long)sub0) * kSubWords};
    const unsigned long long e = reinterpret_cast<const unsigned long long*>(b.entries)[j];
    write_sub(M, g, seg, k, j - seg.sub_off, (uint16_t)(entry_g(e) & kStateMask), b.qpos[j], b.coef, b.dc);
}

} // namespace

cudaError_t launch_huffman(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm)
{
    if (g.num_subs == 0) return cudaSuccess;
    const size_t lut_bytes = (size_t)((g.lut_pool_u16 + 7) / 8) * 16;
    static bool attr_done = false; // per process; all kernels share the same upper bound
    if (!attr_done) {
        const int max_smem = 160 * 1024;
        cudaFuncSetAttribute(huff_sync, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        cudaFuncSetAttribute(huff_link, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        cudaFuncSetAttribute(huff_write, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        attr_done = true;
    }
    const uint32_t chunks = (g.num_subs + kChunkSubs - 1) / kChunkSubs;
    tm->mark(stream, kIdSync);
    huff_sync<<<(chunks + kSyncWarps - 1) / kSyncWarps, kHuffThreads,
                kSyncWinWords * 4 + lut_bytes, stream>>>(g, b);
    tm->mark(stream, kIdLink);
    huff_link<<<1, kLinkThreads, lut_bytes, stream>>>(g, b);
    tm->mark(stream, kIdWrite);
    huff_write<<<(g.num_subs + kHuffThreads - 1) / kHuffThreads, kHuffThreads,
                 kWriteWinWords * 4 + lut_bytes, stream>>>(g, b);
    return cudaGetLastError();
}

} // namespace jb2
