// kernels_destuff.cu — byte unstuffing + restart-segment placement (reference stage a1,
// src/decode_destuff.cu:37-148 + the three CUB scans at :220-316).
//
// Two launches, 16-byte vector loads, no per-byte int arrays:
//   destuff_count : per 4 KiB tile of stuffed bytes, count data bytes and restart markers
//   destuff_apply : tile prefix = sum of the earlier tiles' aggregates (tiny, L2 resident),
//                   block scan, compaction through shared memory, coalesced byte stores.
// Output is the destuffed scan stored as big-endian 32-bit words (byte d lands at d ^ 3) so the
// Huffman kernels read MSB-first words directly; segment k starts at segs[k].sub_off*kSubBytes.
// Algorithmic bytes: Sb read + Sd written (the count pass re-reads Sb from L2).
#include "kernels.hpp"

namespace jb2 {
namespace {

struct Chunk {
    uint32_t keep; // bit i: byte i is a data byte
    uint32_t rst;  // bit i: byte i is the second byte of an RSTn marker
    uint8_t val[kDestuffBytesPerThread]; // emitted value of byte i (0xff for a stuffed 0x00)
};

// The per-byte rule of src/decode_destuff.cu:37-44 and :64-71.
__device__ __forceinline__ void classify(const uint8_t* __restrict__ src, uint32_t n,
                                         uint32_t base, Chunk& ck)
{
    ck.keep = 0;
    ck.rst = 0;
    if (base >= n) return;
    uint32_t w[4];
    if (base + 16 <= n) {
        const uint4 v = *reinterpret_cast<const uint4*>(src + base);
        w[0] = v.x, w[1] = v.y, w[2] = v.z, w[3] = v.w;
    } else {
        w[0] = w[1] = w[2] = w[3] = 0;
        for (uint32_t i = 0; base + i < n; ++i)
            w[i >> 2] |= (uint32_t)src[base + i] << (8 * (i & 3));
    }
    uint32_t prev = base ? src[base - 1] : 0u;
    const uint32_t count = min(16u, n - base);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const uint32_t b = (w[i >> 2] >> (8 * (i & 3))) & 0xffu;
        const bool in = (uint32_t)i < count;
        const bool after_ff = prev == 0xffu;
        const bool keep = in && (after_ff ? b == 0u : b != 0xffu);
        const bool rst = in && after_ff && b >= 0xd0u && b <= 0xd7u;
        ck.keep |= (uint32_t)keep << i;
        ck.rst |= (uint32_t)rst << i;
        ck.val[i] = after_ff ? 0xffu : (uint8_t)b;
        prev = b;
    }
}

// Inclusive block scan of a packed (kept | rst << 16) count; returns the exclusive value.
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* warp_sums,
                                                         uint32_t& block_total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += up;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kDestuffThreads / 32; ++w) {
        const uint32_t s = warp_sums[w];
        if (w < warp) before += s;
        total += s;
    }
    block_total = total;
    return before + inc - v;
}

__global__ void __launch_bounds__(kDestuffThreads)
destuff_count(const uint8_t* __restrict__ src, uint32_t n, unsigned long long* __restrict__ agg)
{
    __shared__ uint32_t warp_sums[kDestuffThreads / 32];
    const uint32_t base = (blockIdx.x * kDestuffThreads + threadIdx.x) * kDestuffBytesPerThread;
    Chunk ck;
    classify(src, n, base, ck);
    uint32_t v = __popc(ck.keep) | (__popc(ck.rst) << 16);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < kDestuffThreads / 32; ++w) t += warp_sums[w];
        agg[blockIdx.x] = (unsigned long long)(t & 0xffffu) | ((unsigned long long)(t >> 16) << 32);
    }
}

__global__ void __launch_bounds__(kDestuffThreads)
destuff_apply(const uint8_t* __restrict__ src, uint32_t n,
              const unsigned long long* __restrict__ agg, const SegDesc* __restrict__ segs,
              int num_segments, uint8_t* __restrict__ dst)
{
    __shared__ uint32_t warp_sums[kDestuffThreads / 32];
    __shared__ unsigned long long red[kDestuffThreads / 32];
    __shared__ uint8_t stage[kDestuffTile];
    __shared__ uint16_t seg_delta[kDestuffTile];

    // tile prefix: data bytes and restart markers in all earlier tiles
    unsigned long long prior = 0;
    for (uint32_t t = threadIdx.x; t < blockIdx.x; t += kDestuffThreads) prior += agg[t];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) prior += __shfl_xor_sync(0xffffffffu, prior, d);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = prior;
    __syncthreads();
    prior = 0;
#pragma unroll
    for (int w = 0; w < kDestuffThreads / 32; ++w) prior += red[w];
    const uint32_t kept_base = (uint32_t)prior;
    const uint32_t seg_base = (uint32_t)(prior >> 32);

    const uint32_t base = (blockIdx.x * kDestuffThreads + threadIdx.x) * kDestuffBytesPerThread;
    Chunk ck;
    classify(src, n, base, ck);
    uint32_t tile_total;
    const uint32_t ex = block_exclusive_scan(__popc(ck.keep) | (__popc(ck.rst) << 16), warp_sums,
                                             tile_total);
    uint32_t pos = ex & 0xffffu, seg = ex >> 16;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if ((ck.keep >> i) & 1u) {
            stage[pos] = ck.val[i];
            seg_delta[pos] = (uint16_t)seg;
            ++pos;
        }
        seg += (ck.rst >> i) & 1u;
    }
    __syncthreads();

    const uint32_t tile_kept = tile_total & 0xffffu;
    // destination of data byte number k of the scan: k + sub_off*kSubBytes - kept_before
    uint32_t adj0 = 0;
    if ((int)seg_base < num_segments)
        adj0 = segs[seg_base].sub_off * (uint32_t)kSubBytes - segs[seg_base].kept_before;
    for (uint32_t o = threadIdx.x; o < tile_kept; o += kDestuffThreads) {
        const uint32_t sd = seg_delta[o];
        uint32_t adj = adj0;
        if (sd) {
            const uint32_t s = seg_base + sd;
            if ((int)s >= num_segments) continue; // cannot happen: host and device rules agree
            adj = segs[s].sub_off * (uint32_t)kSubBytes - segs[s].kept_before;
        } else if ((int)seg_base >= num_segments) {
            continue;
        }
        const uint32_t d = kept_base + o + adj;
        dst[d ^ 3u] = stage[o];
    }
}

} // namespace

cudaError_t launch_destuff(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm)
{
    if (g.scan_bytes == 0) return cudaSuccess;
    const uint32_t tiles = (g.scan_bytes + kDestuffTile - 1) / kDestuffTile;
    tm->mark(stream, kIdDestuffCount);
    destuff_count<<<tiles, kDestuffThreads, 0, stream>>>(b.scan_bytes, g.scan_bytes, b.tile_agg);
    tm->mark(stream, kIdDestuffApply);
    destuff_apply<<<tiles, kDestuffThreads, 0, stream>>>(
        b.scan_bytes, g.scan_bytes, b.tile_agg, b.segs, g.num_segments,
        reinterpret_cast<uint8_t*>(b.destuffed));
    return cudaGetLastError();
}

} // namespace jb2
