// kernels_huffman.cu — self-synchronising parallel Huffman decode (reference stages a4-a11,
// src/decode_huffman.cu:413-880; algorithm family of arXiv 2111.09219).
//
// The reference iterates "flow into the next subsequence until your state equals the stored
// one" with a block-wide reduction per round, and a slow chain overwrites its neighbours'
// states.  Here every subsequence gets an independent guess and the guesses are then verified:
//
//   huff_speculate  thread j cold-starts kRunInSubs subsequences before its own, decodes up to
//                   its boundary to obtain a guessed entry state g_j, then decodes its own
//                   subsequence from g_j giving exit state x_j and slot count n_j.  One 8-byte
//                   record {g, x, n} per subsequence.  No inter-thread communication.
//   huff_repair     thread j checks x_{j-1} == g_j.  On mismatch it re-decodes from x_{j-1} and
//                   keeps flowing forward until its exit state matches the next record's guess.
//                   Records are written with single 8-byte stores, so each is always internally
//                   consistent; if every boundary matches and the segment's first record starts
//                   from the initial state, all records are exact by induction.  Each pass
//                   extends the verified prefix of every segment; later passes exit at once when
//                   the previous pass repaired nothing.
//   huff_fixup      last resort after kRepairPasses dirty passes: one thread per segment walks
//                   its records in order (degenerates to a sequential decode on adversarial
//                   streams, like the reference's TODO at src/decode_huffman.cu:475).
//   sub_scan        exclusive prefix sum of n -> first output slot of every subsequence
//                   (replaces the CUB ExclusiveScanByKey at src/decode_huffman.cu:827-860).
//   huff_write      decode from the verified entry states and scatter non-zero coefficients
//                   into the zeroed DU-major buffer (src/decode_huffman.cu:627-682).
#include "huff_threads.hpp"
#include "kernels.hpp"

namespace jb2 {
namespace {

__constant__ uint8_t c_natural[64] = {JB2_NATURAL_ORDER};

struct SharedTables {
    uint32_t offs[kMaxDuPerMcu]; // dc_off | ac_off << 16
    uint32_t tbls[kMaxDuPerMcu]; // dc table slot | ac table slot << 8
    uint8_t natural[64];
};

extern __shared__ __align__(16) uint16_t s_lut[];

__device__ __forceinline__ void stage_tables(const ScanGeom& g, const uint16_t* __restrict__ pool,
                                             SharedTables& st)
{
    const uint32_t words = (g.lut_pool_u16 + 1) / 2;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(pool);
    uint32_t* dst = reinterpret_cast<uint32_t*>(s_lut);
    for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
    if (threadIdx.x < kMaxDuPerMcu) {
        const int s = threadIdx.x < g.dus_per_mcu ? threadIdx.x : 0;
        st.offs[threadIdx.x] = (uint32_t)g.table_off[g.slot_dc[s]] |
                               ((uint32_t)g.table_off[g.slot_ac[s]] << 16);
        st.tbls[threadIdx.x] = (uint32_t)g.slot_dc[s] | ((uint32_t)g.slot_ac[s] << 8);
    }
    if (threadIdx.x < 64) st.natural[threadIdx.x] = c_natural[threadIdx.x];
    __syncthreads();
}

// Device memory policy of decode_span: LUTs and per-slot tables in shared memory, the destuffed
// scan read word by word through the read-only path (L1 resident).
struct DevMem {
    const SharedTables* st;
    const HuffSlow* slow_;
    __device__ __forceinline__ uint16_t lut(uint32_t i) const { return s_lut[i]; }
    __device__ __forceinline__ uint32_t slot_offs(uint32_t c) const { return st->offs[c]; }
    __device__ __forceinline__ uint32_t slot_tbls(uint32_t c) const { return st->tbls[c]; }
    __device__ __forceinline__ const HuffSlow& slow(uint32_t s) const { return slow_[s]; }
    __device__ __forceinline__ uint32_t natural(uint32_t i) const { return st->natural[i]; }
    __device__ __forceinline__ uint32_t word(const uint32_t* words, uint32_t i) const
    {
        return __ldg(words + i);
    }
    // Predicated load written straight into `dst`, so the scoreboard wait happens where the
    // word is next read (one window slide later), not where it is requested.
    __device__ __forceinline__ void word_if(const uint32_t* words, uint32_t i, bool pred,
                                            uint32_t& dst) const
    {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t"
                     "@p ld.global.nc.u32 %0, [%1];\n\t}"
                     : "+r"(dst)
                     : "l"(words + i), "r"((uint32_t)pred));
    }
};

__global__ void __launch_bounds__(kHuffThreads)
huff_speculate(const ScanGeom g, const ScanBufs b)
{
    __shared__ SharedTables st;
    stage_tables(g, b.lut_pool, st);
    const uint32_t j = blockIdx.x * kHuffThreads + threadIdx.x;
    if (j >= g.num_subs) return;
    const DevMem M{&st, b.slow};
    uint32_t k;
    const unsigned long long e =
        speculate_one(M, make_slot_tables(M, (uint32_t)g.dus_per_mcu), b.segs, g.num_segments, b.destuffed, j, &k);
    b.seg_idx[j] = k;
    reinterpret_cast<unsigned long long*>(b.entries)[j] = e;
}

__global__ void __launch_bounds__(kHuffThreads)
huff_repair(const ScanGeom g, const ScanBufs b, int pass)
{
    __shared__ SharedTables st;
    __shared__ uint32_t s_skip;
    if (threadIdx.x == 0) s_skip = pass > 0 && b.ctrl->repairs[pass - 1] == 0;
    __syncthreads();
    if (s_skip) return;
    stage_tables(g, b.lut_pool, st);
    const uint32_t j = blockIdx.x * kHuffThreads + threadIdx.x;
    if (j >= g.num_subs) return;
    const DevMem M{&st, b.slow};
    volatile unsigned long long* entries = reinterpret_cast<volatile unsigned long long*>(b.entries);
    const uint32_t repaired = repair_one(M, make_slot_tables(M, (uint32_t)g.dus_per_mcu), b.segs, b.seg_idx, b.destuffed, entries, j);
    if (repaired) atomicAdd(&b.ctrl->repairs[pass], repaired);
}

__global__ void __launch_bounds__(32)
huff_fixup(const ScanGeom g, const ScanBufs b)
{
    __shared__ SharedTables st;
    __shared__ uint32_t s_skip;
    if (threadIdx.x == 0) s_skip = b.ctrl->repairs[kRepairPasses - 1] == 0;
    __syncthreads();
    if (s_skip) return;
    stage_tables(g, b.lut_pool, st);
    const uint32_t k = blockIdx.x * 32 + threadIdx.x;
    if (k >= (uint32_t)g.num_segments) return;
    const DevMem M{&st, b.slow};
    fixup_segment(M, make_slot_tables(M, (uint32_t)g.dus_per_mcu), b.segs[k], b.destuffed, reinterpret_cast<unsigned long long*>(b.entries));
}

constexpr int kScanThreads = 1024;

// Single-CTA exclusive prefix sum of entries[].n (N is tens of thousands: one CTA streams it in
// a few microseconds and needs no inter-CTA protocol).  qpos has N + 1 elements.
__global__ void __launch_bounds__(kScanThreads)
sub_scan(const SubEntry* __restrict__ entries, uint32_t n, uint32_t* __restrict__ qpos)
{
    __shared__ uint32_t warp_sums[kScanThreads / 32];
    __shared__ uint32_t s_carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n; base += kScanThreads) {
        const uint32_t idx = base + threadIdx.x;
        const uint32_t v = idx < n ? entries[idx].n : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += up;
        }
        if (lane == 31) warp_sums[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = warp_sums[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, w, d);
                if (lane >= d) w += up;
            }
            warp_sums[lane] = w; // inclusive over warps
        }
        __syncthreads();
        const uint32_t carry = s_carry;
        const uint32_t before = warp ? warp_sums[warp - 1] : 0u;
        if (idx < n) qpos[idx] = carry + before + inc - v;
        __syncthreads();
        if (threadIdx.x == 0) s_carry = carry + warp_sums[kScanThreads / 32 - 1];
        __syncthreads();
    }
    if (threadIdx.x == 0) qpos[n] = s_carry;
}

__global__ void __launch_bounds__(kHuffThreads)
huff_write(const ScanGeom g, const ScanBufs b)
{
    __shared__ SharedTables st;
    stage_tables(g, b.lut_pool, st);
    const uint32_t j = blockIdx.x * kHuffThreads + threadIdx.x;
    if (j >= g.num_subs) return;
    const DevMem M{&st, b.slow};
    write_one(M, make_slot_tables(M, (uint32_t)g.dus_per_mcu), g, b.segs, b.seg_idx, b.destuffed,
              reinterpret_cast<const unsigned long long*>(b.entries), b.qpos, b.coef, j);
}

} // namespace

cudaError_t launch_huffman(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm)
{
    if (g.num_subs == 0) return cudaSuccess;
    const size_t smem = (size_t)((g.lut_pool_u16 + 1) / 2) * 4;
    static bool attr_done = false; // per process; all kernels share the same upper bound
    if (!attr_done) {
        const int max_smem = 96 * 1024;
        cudaFuncSetAttribute(huff_speculate, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        cudaFuncSetAttribute(huff_repair, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        cudaFuncSetAttribute(huff_fixup, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        cudaFuncSetAttribute(huff_write, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        attr_done = true;
    }
    const uint32_t blocks = (g.num_subs + kHuffThreads - 1) / kHuffThreads;
    tm->mark(stream, kIdSpeculate);
    huff_speculate<<<blocks, kHuffThreads, smem, stream>>>(g, b);
    for (int pass = 0; pass < kRepairPasses; ++pass) {
        tm->mark(stream, kIdRepair0 + pass);
        huff_repair<<<blocks, kHuffThreads, smem, stream>>>(g, b, pass);
    }
    tm->mark(stream, kIdFixup);
    huff_fixup<<<(g.num_segments + 31) / 32, 32, smem, stream>>>(g, b);
    tm->mark(stream, kIdSubScan);
    sub_scan<<<1, kScanThreads, 0, stream>>>(b.entries, g.num_subs, b.qpos);
    tm->mark(stream, kIdWrite);
    huff_write<<<blocks, kHuffThreads, smem, stream>>>(g, b);
    return cudaGetLastError();
}

} // namespace jb2
