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

struct __align__(16) SharedTables {
    SlotRec recs[kMaxDuPerMcu]; // device handles: shared-memory byte addresses
    uint32_t tbls[kMaxDuPerMcu];
    uint8_t natural[64];
};

extern __shared__ __align__(16) uint4 s_dyn[]; // [window words | LUT pool]

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
// volatile: keeps the load where it is written (the compiler otherwise sinks the window prefetch
// under the slide predicate, which exposes its latency)
__device__ __forceinline__ uint32_t lds32_pinned(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
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
// barrier in stage()).  lut_addr = shared address of the LUT pool.
__device__ __forceinline__ void make_device_recs(const ScanGeom& g, SharedTables& st, uint32_t lut_addr)
{
    make_slot_recs(g, st.recs, st.tbls);
    const uint32_t recs_addr = smem_addr(st.recs);
    for (int s = 0; s < kMaxDuPerMcu; ++s) {
        SlotRec& r = st.recs[s];
        r.x = lut_addr + r.x * 4u;
        r.y = lut_addr + r.y * 4u;
        r.z = recs_addr + r.z * (uint32_t)sizeof(SlotRec);
        r.w = (lut_addr + (r.w & 0xffffffu) * 4u) | (r.w & 0xff000000u);
    }
}

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
    const uint32_t lut_vec = (g.lut_pool_words + 3) / 4; // pool is padded to 16 B by the planner
    const uint4* lsrc = reinterpret_cast<const uint4*>(b.lut_pool);
    uint4* ldst = s_dyn + WIN_WORDS / 4;
    for (uint32_t i = threadIdx.x; i < lut_vec; i += blockDim.x) ldst[i] = __ldg(lsrc + i);
    if (threadIdx.x == 0) make_device_recs(g, st, smem_addr(ldst));
    if (threadIdx.x >= 32 && threadIdx.x < 96) st.natural[threadIdx.x - 32] = c_natural[threadIdx.x - 32];
    __syncthreads();
}

// Device memory policy of decode_span over the staged window: every handle is a shared-memory
// byte address.  `win_addr` = address of the window word that holds segment-relative word 0.
struct DevMem {
    const SharedTables* st;
    const HuffSlow* slow_;
    uint32_t recs_addr;
    uint32_t win_addr;
    static constexpr uint32_t kWordStep = 4;
    __device__ __forceinline__ uint32_t lut(uint32_t tbl, uint32_t i) const { return lds32(tbl + i * 4u); }
    __device__ __forceinline__ uint32_t slot_handle(uint32_t c) const
    {
        return recs_addr + c * (uint32_t)sizeof(SlotRec);
    }
    __device__ __forceinline__ SlotRec rec(uint32_t h) const { return lds_rec(h); }
    __device__ __forceinline__ uint32_t slot_tbls(uint32_t c) const { return st->tbls[c]; }
    __device__ __forceinline__ const HuffSlow& slow(uint32_t s) const { return slow_[s]; }
    __device__ __forceinline__ uint32_t natural(uint32_t i) const { return st->natural[i]; }
    __device__ __forceinline__ uint32_t word_handle(uint32_t i) const { return win_addr + i * 4u; }
    __device__ __forceinline__ uint32_t word(uint32_t h) const { return lds32_pinned(h); }
    __device__ __forceinline__ void note_long_code() const {}
};

// Fallback policy of huff_link: tables in shared memory (LUT pool at s_dyn), the scan read from
// global memory.
struct GlobalMem {
    const SharedTables* st;
    const HuffSlow* slow_;
    uint32_t recs_addr;
    const uint32_t* words;
    static constexpr uint32_t kWordStep = 1;
    __device__ __forceinline__ uint32_t lut(uint32_t tbl, uint32_t i) const { return lds32(tbl + i * 4u); }
    __device__ __forceinline__ uint32_t slot_handle(uint32_t c) const
    {
        return recs_addr + c * (uint32_t)sizeof(SlotRec);
    }
    __device__ __forceinline__ SlotRec rec(uint32_t h) const { return lds_rec(h); }
    __device__ __forceinline__ uint32_t slot_tbls(uint32_t c) const { return st->tbls[c]; }
    __device__ __forceinline__ const HuffSlow& slow(uint32_t s) const { return slow_[s]; }
    __device__ __forceinline__ uint32_t natural(uint32_t i) const { return st->natural[i]; }
    __device__ __forceinline__ uint32_t word_handle(uint32_t i) const { return i; }
    __device__ __forceinline__ uint32_t word(uint32_t h) const { return __ldg(words + h); }
    __device__ __forceinline__ void note_long_code() const {}
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

__global__ void __launch_bounds__(kHuffThreads, 1)
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
    DevMem M{&st, b.slow, smem_addr(st.recs), 0};
    if (active) {
        k = find_segment(b.segs, g.num_segments, (uint32_t)sub);
        const SegDesc seg = b.segs[k];
        i = (uint32_t)sub - seg.sub_off;
        head = i == 0;
        seg_bits = seg.nbytes * 8u;
        M.win_addr = smem_addr(s_dyn) + (uint32_t)((int)((long long)seg.sub_off - win_sub0) * kSubBytes);
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

__global__ void __launch_bounds__(kLinkThreads, 1)
huff_link(const ScanGeom g, const ScanBufs b)
{
    __shared__ SharedTables st;
    __shared__ uint32_t s_sum[kLinkThreads / 32], s_flag[kLinkThreads / 32];
    volatile unsigned long long* entries = reinterpret_cast<volatile unsigned long long*>(b.entries);
    const uint32_t N = g.num_subs;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long t_start = 0, t_verified = 0;
    if (threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));

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
            const uint32_t lut_vec = (g.lut_pool_words + 3) / 4;
            for (uint32_t v = threadIdx.x; v < lut_vec; v += kLinkThreads)
                s_dyn[v] = __ldg(reinterpret_cast<const uint4*>(b.lut_pool) + v);
            if (threadIdx.x == 0) make_device_recs(g, st, smem_addr(s_dyn));
            if (threadIdx.x < 64) st.natural[threadIdx.x] = c_natural[threadIdx.x];
            staged = true;
            __syncthreads();
        }
        uint32_t rewritten = 0;
        for (uint32_t j = stride * (1u + threadIdx.x); j < N; j += stride * kLinkThreads) {
            if (boundary_ok(entries[j - 1], entries[j])) continue;
            const SegDesc seg = b.segs[b.seg_idx[j]];
            const GlobalMem M{&st, b.slow, smem_addr(st.recs),
                              b.destuffed + (size_t)seg.sub_off * kSubWords};
            rewritten += relink_from(M, seg, entries, j);
        }
        slow_total += rewritten;
        ++rounds;
        stride = 1; // concurrent repairs may interleave: from now on check every boundary
        __syncthreads();
    }

    if (threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_verified));
    // ---- segmented exclusive scan of the slot counts ---------------------------------------
    // Tiles of 4 records per thread: coalesced 32-byte loads and 16-byte stores (a strided
    // per-thread run made every load its own sector and took 10x longer on one SM).
    __shared__ uint32_t s_carry, s_excl[kLinkThreads / 32];
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const unsigned long long* ent = reinterpret_cast<const unsigned long long*>(b.entries);
    constexpr uint32_t kPer = 4, kTile = kLinkThreads * kPer;
    for (uint32_t base = 0; base < N; base += kTile) {
        const uint32_t j0 = base + threadIdx.x * kPer;
        unsigned long long e[kPer];
        if (j0 + kPer <= N) {
            const ulonglong2 a0 = __ldcg(reinterpret_cast<const ulonglong2*>(ent + j0));
            const ulonglong2 a1 = __ldcg(reinterpret_cast<const ulonglong2*>(ent + j0 + 2));
            e[0] = a0.x, e[1] = a0.y, e[2] = a1.x, e[3] = a1.y;
        } else {
#pragma unroll
            for (uint32_t i = 0; i < kPer; ++i) e[i] = j0 + i < N ? __ldcg(ent + j0 + i) : 0ull;
        }
        uint32_t sum = 0, flag = 0;
#pragma unroll
        for (uint32_t i = 0; i < kPer; ++i) {
            if (entry_g(e[i]) & kStateHead) sum = 0, flag = 1;
            sum += entry_n(e[i]);
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
        if (warp == 0) { // second level: segmented scan of the 32 warp aggregates
            uint32_t v = s_sum[lane], gf = s_flag[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, v, d);
                const uint32_t upg = __shfl_up_sync(0xffffffffu, gf, d);
                if (lane >= d) {
                    if (!gf) v += up;
                    gf |= upg;
                }
            }
            const uint32_t tile_carry = s_carry;
            const uint32_t incl = gf ? v : v + tile_carry; // running value after warp `lane`
            uint32_t excl = __shfl_up_sync(0xffffffffu, incl, 1);
            if (lane == 0) excl = tile_carry;
            s_excl[lane] = excl;
            if (lane == 31) s_carry = incl; // every lane read tile_carry before the shuffles
        }
        __syncthreads();
        // running value before this thread's records
        uint32_t acc = __shfl_up_sync(0xffffffffu, inc, 1);
        const uint32_t fex = __shfl_up_sync(0xffffffffu, f, 1);
        if (lane == 0) acc = s_excl[warp];
        else if (!fex) acc += s_excl[warp];
        uint32_t out[kPer];
#pragma unroll
        for (uint32_t i = 0; i < kPer; ++i) {
            if (entry_g(e[i]) & kStateHead) acc = 0;
            out[i] = acc;
            acc += entry_n(e[i]);
        }
        if (j0 + kPer <= N) {
            *reinterpret_cast<uint4*>(b.qpos + j0) = make_uint4(out[0], out[1], out[2], out[3]);
        } else {
#pragma unroll
            for (uint32_t i = 0; i < kPer; ++i)
                if (j0 + i < N) b.qpos[j0 + i] = out[i];
        }
        // no barrier here: s_sum / s_excl are rewritten only behind the next tile's barriers
    }
    if (slow_total) atomicAdd(&b.ctrl->slow_repairs, slow_total);
    if (threadIdx.x == 0) {
        unsigned long long t_end;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
        b.ctrl->link_rounds = rounds;
        b.ctrl->link_verify_ns = (uint32_t)(t_verified - t_start);
        b.ctrl->link_scan_ns = (uint32_t)(t_end - t_verified);
    }
}

__global__ void __launch_bounds__(kHuffThreads, 1)
huff_write(const ScanGeom g, const ScanBufs b)
{
    __shared__ SharedTables st;
    const uint32_t sub0 = blockIdx.x * kHuffThreads;
    stage<kWriteWinWords>(g, b, st, (long long)sub0 * kSubWords);
    const uint32_t j = sub0 + threadIdx.x;
    if (j >= g.num_subs) return;
    const uint32_t k = b.seg_idx[j];
    const SegDesc seg = b.segs[k];
    const DevMem M{&st, b.slow, smem_addr(st.recs),
                   smem_addr(s_dyn) + (uint32_t)((int)((long long)seg.sub_off - (long long)sub0) * kSubBytes)};
    const unsigned long long e = reinterpret_cast<const unsigned long long*>(b.entries)[j];
    write_sub(M, g, seg, k, j - seg.sub_off, (uint16_t)(entry_g(e) & kStateMask), b.qpos[j], b.coef, b.dc);
}

} // namespace

cudaError_t launch_huffman(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm)
{
    if (g.num_subs == 0) return cudaSuccess;
    const size_t lut_bytes = (size_t)((g.lut_pool_words + 3) / 4) * 16;
    static std::atomic<unsigned long long> configured{0};
    if (first_use_on_device(configured)) {
        const int max_smem = 200 * 1024;
        prefer_max_shared(huff_sync, max_smem);
        prefer_max_shared(huff_link, max_smem);
        prefer_max_shared(huff_write, max_smem);
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
