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
//   huff_write  decode from the verified entry states, two threads per subsequence (the record
//               holds the state at its middle).  A warp assembles the data units that lie
//               entirely inside its slots in a shared-memory buffer and flushes them with
//               coalesced 16-byte stores; data units shared with a neighbouring warp are scattered
//               into the zero-filled global buffer.  Coefficients keep their zig-zag index (the
//               IDCT un-zig-zags in registers); DC differences also go to a compact per-DU array
//               (src/decode_huffman.cu:627-682).
#include "huff_threads.hpp"
#include "kernels.hpp"

namespace jb2 {
namespace {

constexpr int kSyncWarps = kHuffThreads / 32;
constexpr int kSyncUseful = kSyncWarps * kChunkSubs;            // records written per CTA
constexpr int kSyncWinSubs = kSyncUseful + kOverlapSubs;        // subsequences staged per CTA
constexpr int kWinMargin = 4;                                   // words past the last subsequence
constexpr int kSyncWinWords = kSyncWinSubs * kSubWords + kWinMargin;
constexpr int kWriteThreads = 320;  // at most; two threads per subsequence
constexpr int kTileDus = 32;        // data units a warp assembles at a time
// non-zero coefficients one thread can meet in half a subsequence: each takes >= 2 bits, plus the
// symbol that straddles the end
constexpr int kListCap = kSubBits / 4 + 2;
constexpr int kListStride = kListCap | 1;                          // words between the lanes' lists
constexpr int kWarpListBytes = (32 * kListStride * 4 + 15) & ~15;
#if defined(JB2_WRITE_TILES)
constexpr int kWarpSmem = kWarpListBytes + kTileDus * 128;         // lists + tile of one warp
#else
constexpr int kWarpSmem = kWarpListBytes;
#endif
static_assert(kSubBytes <= 128, "slot offsets inside half a subsequence must fit 16 bits");
static_assert(kSyncWinWords % 4 == 0 && kSubWords % 4 == 0, "window must be uint4 sized");

struct __align__(16) SharedTables {
    SlotRec recs[kMaxDuPerMcu]; // device handles: shared-memory byte addresses
    uint32_t tbls[kMaxDuPerMcu];
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
__device__ __forceinline__ void stage(const ScanGeom& g, const ScanBufs& b, SharedTables& st,
                                      long long first_word, const int WIN_WORDS)
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
    __device__ __forceinline__ uint32_t word_handle(uint32_t i) const { return win_addr + i * 4u; }
    __device__ __forceinline__ uint32_t word(uint32_t h) const { return lds32_pinned(h); }
    __device__ __forceinline__ void note_long_code() const {}
    __device__ __forceinline__ void put(uint32_t, int) const {}
};

// Policy of huff_write: DevMem + the coefficient sink.  The symbol loop only appends
// {slot - slot0, value} to the thread's list in shared memory (one contiguous region per lane, an
// odd number of words apart); huff_write turns the lists into data units afterwards, all lanes
// working on one list at a time.
struct WriteMem : DevMem {
    mutable uint32_t list_ptr; // shared address of the next entry
    uint32_t slot0;            // first slot of this thread's half subsequence (segment relative)
    __device__ __forceinline__ void put(uint32_t slot, int v) const
    {
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(list_ptr),
                     "r"(((slot - slot0) << 16) | ((uint32_t)v & 0xffffu)));
        list_ptr += 4u;
    }
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
    __device__ __forceinline__ uint32_t word_handle(uint32_t i) const { return i; }
    __device__ __forceinline__ uint32_t word(uint32_t h) const { return __ldg(words + h); }
    __device__ __forceinline__ void note_long_code() const {}
    __device__ __forceinline__ void put(uint32_t, int) const {}
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
    const unsigned long long t0 = now_ns();
    const long long win_sub0 = (long long)blockIdx.x * kSyncUseful - kOverlapSubs;
    stage(g, b, st, win_sub0 * kSubWords, kSyncWinWords);
    const unsigned long long t1 = now_ns();
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
    uint32_t mid = 0;
    unsigned long long rec = active ? decode_sub(M, seg_bits, i, entry, &mid) : 0ull;
    auto iterate = [&](int frozen) {
        for (;;) {
            const uint16_t px = (uint16_t)__shfl_up_sync(0xffffffffu, (uint32_t)entry_x(rec), 1);
            const uint16_t want = (lane <= frozen || head || !active) ? entry : px;
            const bool changed = want != entry;
            if (!__any_sync(0xffffffffu, changed)) break;
            if (changed) {
                entry = want;
                rec = decode_sub(M, seg_bits, i, entry, &mid);
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
                rec = decode_sub(M, seg_bits, i, entry, &mid);
                atomicAdd(&b.ctrl->relinked, 1u);
            }
            iterate(kOverlapSubs);
            if (lane == last_lane) mailbox[chunk] = 0x80000000u | entry_x(rec);
        }
    }
    if (lane == 0) {
        atomicMax(&b.ctrl->sync_stage_ns, (uint32_t)(t1 - t0));
        atomicMax(&b.ctrl->sync_iter_ns, (uint32_t)(now_ns() - t1));
    }
    if (active && lane >= kOverlapSubs) {
        reinterpret_cast<ulonglong2*>(b.entries)[sub] = make_ulonglong2(
            pack_entry((uint16_t)(entry | (head ? kStateHead : 0)), entry_x(rec), entry_n(rec)),
            (unsigned long long)mid);
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
            bad |= boundary_ok(entries[2 * (size_t)j - 2], entries[2 * (size_t)j]) ? 0u : 1u;
        if (!__syncthreads_or((int)bad)) break;
        if (!staged) { // tables are only needed on this path
            const uint32_t lut_vec = (g.lut_pool_words + 3) / 4;
            for (uint32_t v = threadIdx.x; v < lut_vec; v += kLinkThreads)
                s_dyn[v] = __ldg(reinterpret_cast<const uint4*>(b.lut_pool) + v);
            if (threadIdx.x == 0) make_device_recs(g, st, smem_addr(s_dyn));
            staged = true;
            __syncthreads();
        }
        uint32_t rewritten = 0;
        for (uint32_t j = stride * (1u + threadIdx.x); j < N; j += stride * kLinkThreads) {
            if (boundary_ok(entries[2 * (size_t)j - 2], entries[2 * (size_t)j])) continue;
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
    // Tiles of 4 records per thread: 64 contiguous bytes loaded and 16 bytes stored per thread
    // (a strided per-thread run made every load its own sector and took 10x longer on one SM).
    __shared__ uint32_t s_carry, s_excl[kLinkThreads / 32];
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const unsigned long long* ent = reinterpret_cast<const unsigned long long*>(b.entries);
    constexpr uint32_t kPer = 4, kTile = kLinkThreads * kPer;
    for (uint32_t base = 0; base < N; base += kTile) {
        const uint32_t j0 = base + threadIdx.x * kPer;
        unsigned long long e[kPer];
        if (j0 + kPer <= N) {
            const ulonglong2* rp = reinterpret_cast<const ulonglong2*>(ent) + j0;
#pragma unroll
            for (uint32_t i = 0; i < kPer; ++i) e[i] = __ldcg(rp + i).x;
        } else {
#pragma unroll
            for (uint32_t i = 0; i < kPer; ++i) e[i] = j0 + i < N ? __ldcg(ent + 2 * (size_t)(j0 + i)) : 0ull;
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

__global__ void __launch_bounds__(kWriteThreads, 1)
huff_write(const ScanGeom g, const ScanBufs b)
{
    __shared__ SharedTables st;
    const unsigned long long t0 = now_ns();
    const uint32_t subs_per_cta = blockDim.x / 2;
    const int win_words = (int)(subs_per_cta * kSubWords + kWinMargin);
    const uint32_t sub0 = blockIdx.x * subs_per_cta;
    stage(g, b, st, (long long)sub0 * kSubWords, win_words);
    const unsigned long long t1 = now_ns();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t j = sub0 + (threadIdx.x >> 1), h = threadIdx.x & 1u;

    // this thread's half subsequence: record, placement, slot range [qs, qe)
    bool has = false;
    unsigned long long rec = 0;
    uint32_t mid = 0, q0 = 0, first_du = 0, q_limit = 0, qs = 0, qe = 0, i = 0;
    SegDesc seg = {0, 0, 0, 0};
    if (j < g.num_subs) {
        const ulonglong2 r = __ldg(reinterpret_cast<const ulonglong2*>(b.entries) + j);
        rec = r.x;
        mid = (uint32_t)r.y;
        const uint32_t k = b.seg_idx[j];
        seg = b.segs[k];
        i = j - seg.sub_off;
        q0 = b.qpos[j];
        if (segment_extent(g, k, first_du, q_limit)) {
            qs = h ? q0 + (mid >> 16) : q0;
            qe = umin(h ? q0 + entry_n(rec) : q0 + (mid >> 16), q_limit);
            has = qe > qs;
        }
    }
    const uint32_t lut_vec = (g.lut_pool_words + 3) / 4;
    uint8_t* warp_mem = reinterpret_cast<uint8_t*>(s_dyn + win_words / 4 + lut_vec) + (size_t)warp * kWarpSmem;
    const uint32_t list_addr = smem_addr(warp_mem) + (uint32_t)lane * (kListStride * 4u);
    uint32_t list_end = list_addr;
    if (has) {
        WriteMem M;
        M.st = &st;
        M.slow_ = b.slow;
        M.recs_addr = smem_addr(st.recs);
        M.win_addr = smem_addr(s_dyn) + (uint32_t)((int)((long long)seg.sub_off - (long long)sub0) * kSubBytes);
        M.list_ptr = list_addr;
        M.slot0 = qs;
        write_half(M, seg, i, h, rec, mid, q0, q_limit);
        list_end = M.list_ptr;
    }
    __syncwarp();
    const unsigned long long t2 = now_ns();

    // ---- lists -> data units --------------------------------------------------------------
    // The warp's slots [gmin, gmax) are contiguous up to gaps at restart boundaries.  Data units
    // that lie entirely inside belong to this warp alone: they are assembled in the tile (zeros
    // included) and stored with coalesced 16-byte stores.  The first / last data unit may be
    // shared with the neighbouring warp: only their non-zero coefficients are stored, into the
    // buffer huff_sync zero-filled.
    const uint32_t gs0 = first_du * 64u + qs; // global slot of this thread's first coefficient
    const uint32_t gmin = __reduce_min_sync(0xffffffffu, has ? gs0 : 0xffffffffu);
    const uint32_t gmax = __reduce_max_sync(0xffffffffu, has ? first_du * 64u + qe : 0u);
    uint32_t tiles = 0;
#if !defined(JB2_WRITE_TILES)
    // Every thread scatters its own list straight into the zero-filled buffers.  (The tile
    // assembly below stores coalesced 16-byte vectors instead, but its ~3000 extra warp
    // instructions per warp and its shared-memory tiles (one wave fewer CTAs per SM) made the
    // kernel 15% slower on B200 than these 2-byte stores into L2-resident lines.)
    for (uint32_t at = list_addr; at != list_end; at += 4u) {
        const uint32_t e = lds32(at);
        const uint32_t slot = gs0 + (e >> 16);
        if ((slot & 63u) == 0) b.dc[slot >> 6] = (int16_t)(e & 0xffffu);
        else b.coef[slot] = (int16_t)(e & 0xffffu);
    }
    if (false) {
#else
    if (gmax > 0) {
#endif
        const uint32_t own_lo = (gmin + 63u) >> 6, own_hi = gmax >> 6; // owned: [own_lo, own_hi)
        uint16_t* tile = reinterpret_cast<uint16_t*>(warp_mem + kWarpListBytes);
        uint4* tile_vec = reinterpret_cast<uint4*>(tile);
        uint32_t rd = list_addr;
        for (;; ++tiles) {
            // next tile = the one holding the smallest pending slot (empty stretches of the
            // warp's range, e.g. flat image areas, stay as huff_sync zero-filled them)
            const uint32_t pending = rd != list_end ? gs0 + (lds32(rd) >> 16) : 0xffffffffu;
            const uint32_t first = __reduce_min_sync(0xffffffffu, pending);
            if (first == 0xffffffffu) break;
            const uint32_t du0 = first >> 6;
#pragma unroll
            for (int v = 0; v < kTileDus * 8 / 32; ++v) tile_vec[v * 32 + lane] = make_uint4(0u, 0u, 0u, 0u);
            __syncwarp();
            const uint32_t tile_slot0 = du0 * 64u, tile_slots = kTileDus * 64u;
            // lists with entries in this tile, one list at a time, a batch of 32 entries per step
            // (a lane walking its own list alone would serialise the whole tile behind the lane
            // with the longest list)
            uint32_t todo = __ballot_sync(0xffffffffu, pending - tile_slot0 < tile_slots);
            while (todo) {
                const int src = __ffs((int)todo) - 1;
                todo &= todo - 1u;
                const uint32_t src_rd = __shfl_sync(0xffffffffu, rd, src);
                const uint32_t src_end = __shfl_sync(0xffffffffu, list_end, src);
                const uint32_t src_gs0 = __shfl_sync(0xffffffffu, gs0, src);
                uint32_t consumed = 0;
                for (uint32_t at = src_rd + (uint32_t)lane * 4u;; at += 128u) {
                    bool in = false;
                    if (at < src_end) {
                        const uint32_t e = lds32(at);
                        const uint32_t rel = src_gs0 + (e >> 16) - tile_slot0;
                        in = rel < tile_slots; // entries are sorted: the hits form a prefix
                        if (in) tile[rel] = (uint16_t)e;
                    }
                    const uint32_t hits = __popc(__ballot_sync(0xffffffffu, in));
                    consumed += hits;
                    if (hits < 32u) break;
                }
                if (lane == src) rd += consumed * 4u;
            }
            __syncwarp();
#pragma unroll
            for (int v = 0; v < kTileDus * 8 / 32; ++v) {
                const uint32_t idx = (uint32_t)v * 32u + (uint32_t)lane; // 16-byte vector of the tile
                const uint32_t du = du0 + (idx >> 3);
                const uint4 val = tile_vec[idx];
                // eight consecutive lanes hold one data unit: all-zero data units are not stored
                const uint32_t nz = __ballot_sync(0xffffffffu, (val.x | val.y | val.z | val.w) != 0);
                if (((nz >> (lane & 24)) & 0xffu) == 0) continue;
                if (du >= own_lo && du < own_hi) {
                    reinterpret_cast<uint4*>(b.coef)[(size_t)du * 8 + (idx & 7u)] = val;
                    if ((idx & 7u) == 0) b.dc[du] = (int16_t)(val.x & 0xffffu); // input of the DC scan
                } else if (du * 64u < gmax && du >= (gmin >> 6)) { // shared data unit
                    const uint32_t w[4] = {val.x, val.y, val.z, val.w};
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const uint16_t c16 = (uint16_t)(w[k >> 1] >> (16 * (k & 1)));
                        const uint32_t slot = du * 64u + (idx & 7u) * 8u + (uint32_t)k;
                        if (c16 != 0) {
                            if ((slot & 63u) == 0) b.dc[du] = (int16_t)c16;
                            else b.coef[slot] = (int16_t)c16;
                        }
                    }
                }
            }
            __syncwarp();
        }
    }
    if (lane == 0) {
        atomicMax(&b.ctrl->write_stage_ns, (uint32_t)(t1 - t0));
        atomicMax(&b.ctrl->write_decode_ns, (uint32_t)(t2 - t1));
        atomicMax(&b.ctrl->write_flush_ns, (uint32_t)(now_ns() - t2));
        atomicMax(&b.ctrl->write_tiles_max, tiles);
        atomicAdd(&b.ctrl->write_tiles, tiles);
    }
}

} // namespace

cudaError_t launch_huffman(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm)
{
    if (g.num_subs == 0) return cudaSuccess;
    const size_t lut_bytes = (size_t)((g.lut_pool_words + 3) / 4) * 16;
    static std::atomic<unsigned long long> configured{0};
    if (first_use_on_device(configured)) {
        const int max_smem = 226 * 1024; // 227 KiB per CTA minus the static part
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
    // as many warps per CTA as the shared memory left by the LUT pool allows (each warp needs
    // its coefficient lists and one tile)
    const size_t budget = 225 * 1024;
    int warps = kWriteThreads / 32;
    auto write_smem = [&](int w) {
        return (size_t)(w * 16 * kSubWords + kWinMargin) * 4 + lut_bytes + (size_t)w * kWarpSmem;
    };
    while (warps > 1 && write_smem(warps) + 1024 > budget) --warps;
    const uint32_t subs_per_cta = (uint32_t)warps * 16u;
    tm->mark(stream, kIdWrite);
    huff_write<<<(g.num_subs + subs_per_cta - 1) / subs_per_cta, warps * 32, write_smem(warps), stream>>>(g, b);
    return cudaGetLastError();
}

} // namespace jb2
