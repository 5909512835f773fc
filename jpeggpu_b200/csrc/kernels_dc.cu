// kernels_dc.cu — DC un-differencing as a device-wide segmented prefix sum (reference stage a12,
// src/decode_dc.cu:42-169: one CUB InclusiveSum[ByKey] per component over a 128-byte-strided
// permutation iterator).  Here: one pass for all components of the scan.
//
// Work unit = MCU.  Restart boundaries fall on multiples of the restart interval, so the carry
// into a tile of kDcTile MCUs is pure arithmetic over per-tile aggregates:
// Input: the compact array of DC differences huff_write produced (one int16 per DU, coded
// order: contiguous, unlike the reference's 128-byte-strided permutation iterator); output: the
// absolute DC value stored into slot 0 of every DU of the coefficient buffer.
//   dc_tile_sums : agg[t][a] = sum of component a's DC differences over the MCUs of tile t that
//                  follow the last restart boundary inside t (whole tile if none)
//   dc_apply     : carry = sum of agg over the tiles back to the one holding the segment start,
//                  then a segmented block scan and the in-place write of absolute DC values.
// Arithmetic is int32 truncated to int16 on store == the reference's int16 wrap-around sums.
// Algorithmic bytes: 4 per DU (read + write one int16); touched sectors are 32 B each.
#include "kernels.hpp"

namespace jb2 {
namespace {

struct Dc4 {
    int v[kMaxComp];
};

__device__ __forceinline__ void load_mcu(const ScanGeom& g, const int16_t* __restrict__ dc,
                                         uint32_t mcu, int (&diff)[kMaxDuPerMcu], Dc4& tot)
{
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a) tot.v[a] = 0;
    const int16_t* base = dc + (size_t)mcu * g.dus_per_mcu;
#pragma unroll
    for (int s = 0; s < kMaxDuPerMcu; ++s) {
        diff[s] = 0;
        if (s < g.dus_per_mcu) {
            diff[s] = base[s];
            const int a = g.slot_comp[s];
#pragma unroll
            for (int k = 0; k < kMaxComp; ++k)
                if (k == a) tot.v[k] += diff[s];
        }
    }
}

// Segmented inclusive scan over the block: `head` marks the first MCU of a restart segment.
// Returns the inclusive value; *tail_sum receives the block aggregate after the last head.
__device__ __forceinline__ Dc4 block_segmented_scan(Dc4 val, bool head, Dc4* warp_val,
                                                    uint32_t* warp_flag, Dc4& tail_sum)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t flag = head ? 1u : 0u;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Dc4 up;
#pragma unroll
        for (int a = 0; a < kMaxComp; ++a) up.v[a] = __shfl_up_sync(0xffffffffu, val.v[a], d);
        const uint32_t upf = __shfl_up_sync(0xffffffffu, flag, d);
        if (lane >= d) {
            if (!flag) {
#pragma unroll
                for (int a = 0; a < kMaxComp; ++a) val.v[a] += up.v[a];
            }
            flag |= upf;
        }
    }
    if (lane == 31) {
        warp_val[warp] = val;
        warp_flag[warp] = flag;
    }
    __syncthreads();
    // carry from earlier warps: walk back until a warp that contains a head
    Dc4 carry;
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a) carry.v[a] = 0;
    if (!flag) {
        for (int w = warp - 1; w >= 0; --w) {
#pragma unroll
            for (int a = 0; a < kMaxComp; ++a) carry.v[a] += warp_val[w].v[a];
            if (warp_flag[w]) break;
        }
    }
    // block aggregate after the last head
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a) tail_sum.v[a] = 0;
    for (int w = kDcTile / 32 - 1; w >= 0; --w) {
#pragma unroll
        for (int a = 0; a < kMaxComp; ++a) tail_sum.v[a] += warp_val[w].v[a];
        if (warp_flag[w]) break;
    }
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a) val.v[a] += carry.v[a];
    return val;
}

__global__ void __launch_bounds__(kDcTile)
dc_tile_sums(const ScanGeom g, const int16_t* __restrict__ dc, int* __restrict__ agg)
{
    __shared__ Dc4 warp_val[kDcTile / 32];
    __shared__ uint32_t warp_flag[kDcTile / 32];
    const uint32_t mcu = blockIdx.x * kDcTile + threadIdx.x;
    const bool valid = mcu < (uint32_t)g.total_mcus;
    int diff[kMaxDuPerMcu];
    Dc4 tot;
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a) tot.v[a] = 0;
    if (valid) load_mcu(g, dc, mcu, diff, tot);
    const bool head = valid && (mcu % (uint32_t)g.restart_interval) == 0;
    Dc4 tail;
    block_segmented_scan(tot, head, warp_val, warp_flag, tail);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int a = 0; a < kMaxComp; ++a) agg[blockIdx.x * kMaxComp + a] = tail.v[a];
    }
}

__global__ void __launch_bounds__(kDcTile)
dc_apply(const ScanGeom g, const int16_t* __restrict__ dc, int16_t* __restrict__ coef,
         const int* __restrict__ agg)
{
    __shared__ Dc4 warp_val[kDcTile / 32];
    __shared__ uint32_t warp_flag[kDcTile / 32];
    __shared__ Dc4 s_red[kDcTile / 32];
    const uint32_t tile_mcu0 = blockIdx.x * kDcTile;
    const uint32_t ri = (uint32_t)g.restart_interval;
    // tiles [first_tile, blockIdx.x) carry into this one
    const uint32_t seg_start = tile_mcu0 / ri * ri;
    const uint32_t first_tile = seg_start / kDcTile;
    Dc4 carry;
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a) carry.v[a] = 0;
    for (uint32_t t = first_tile + threadIdx.x; t < blockIdx.x; t += kDcTile) {
#pragma unroll
        for (int a = 0; a < kMaxComp; ++a) carry.v[a] += agg[t * kMaxComp + a];
    }
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) carry.v[a] += __shfl_xor_sync(0xffffffffu, carry.v[a], d);
    }
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = carry;
    __syncthreads();
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a) {
        carry.v[a] = 0;
        for (int w = 0; w < kDcTile / 32; ++w) carry.v[a] += s_red[w].v[a];
    }

    const uint32_t mcu = tile_mcu0 + threadIdx.x;
    const bool valid = mcu < (uint32_t)g.total_mcus;
    int diff[kMaxDuPerMcu];
    Dc4 tot;
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a) tot.v[a] = 0;
    if (valid) load_mcu(g, dc, mcu, diff, tot);
    const bool head = valid && (mcu % ri) == 0;
    Dc4 tail;
    const Dc4 inc = block_segmented_scan(tot, head, warp_val, warp_flag, tail);
    if (!valid) return;
    // the tile carry applies only while no restart boundary has been seen inside the tile
    const bool fresh = (mcu / ri * ri) >= tile_mcu0;
    Dc4 pred; // predictor before this MCU
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a)
        pred.v[a] = inc.v[a] - tot.v[a] + (fresh ? 0 : carry.v[a]);
    int16_t* base = coef + (size_t)mcu * g.dus_per_mcu * 64;
#pragma unroll
    for (int s = 0; s < kMaxDuPerMcu; ++s) {
        if (s < g.dus_per_mcu) {
            const int a = g.slot_comp[s];
            int value = 0;
#pragma unroll
            for (int k = 0; k < kMaxComp; ++k)
                if (k == a) {
                    pred.v[k] += diff[s];
                    value = pred.v[k];
                }
            base[s * 64] = (int16_t)value;
        }
    }
}

} // namespace

cudaError_t launch_dc(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm)
{
    if (g.total_mcus == 0) return cudaSuccess;
    static std::atomic<unsigned long long> configured{0};
    if (first_use_on_device(configured)) {
        prefer_max_shared(dc_tile_sums);
        prefer_max_shared(dc_apply);
    }
    const uint32_t tiles = ((uint32_t)g.total_mcus + kDcTile - 1) / kDcTile;
    tm->mark(stream, kIdDcSums);
    dc_tile_sums<<<tiles, kDcTile, 0, stream>>>(g, b.dc, b.dc_agg);
    tm->mark(stream, kIdDcApply);
    dc_apply<<<tiles, kDcTile, 0, stream>>>(g, b.dc, b.coef, b.dc_agg);
    return cudaGetLastError();
}

} // namespace jb2
