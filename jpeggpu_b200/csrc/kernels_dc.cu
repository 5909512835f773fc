// kernels_dc.cu — DC un-differencing as a device-wide segmented prefix sum (reference stage a12,
// src/decode_dc.cu:42-169: one CUB InclusiveSum[ByKey] per component over a 128-byte-strided
// permutation iterator).  Here: one pass for all components of the scan.
//
// Work unit = MCU.  Restart boundaries fall on multiples of the restart interval, so the carry
// into a tile of kDcTile MCUs is pure arithmetic over per-tile aggregates:
// Input: the compact array of DC differences huff_write produced (one int16 per DU, coded
// order: contiguous, unlike the reference's 128-byte-strided permutation iterator); output: the
// absolute DC values in the same array.
//   agg[t][a] = sum of component a's DC differences over the MCUs of tile t that follow the last
//               restart boundary inside t (whole tile if none)
//   carry     = sum of agg over the tiles back to the one holding the segment start; then the
//               segmented block scan result is written back IN PLACE into the compact array
//               (the IDCT / transpose pick the DC from there instead of coefficient slot 0).
// Arithmetic is int32 truncated to int16 on store == the reference's int16 wrap-around sums.
// Algorithmic bytes: 4 per DU (read + write one int16); touched sectors are 32 B each.
#include "kernels.hpp"

#include <cstddef>

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

// Tile aggregates travel in two 64-bit words, each {valid:1 | sum of component 2k+1:16 | sum of
// component 2k:16}: DC arithmetic is modulo 2^16 (the reference accumulates in int16), so the
// payload fits the flagged word itself and neither a fence nor a second load is needed.
__device__ __forceinline__ unsigned long long ld_relaxed64(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed64(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
constexpr unsigned long long kDcValid = 1ull << 32;

// One launch: every CTA scans its tile, publishes the tile aggregate (sum after the tile's last
// restart boundary) and then sums the aggregates of the tiles between the start of its restart
// segment and itself.  Aggregates do not depend on other tiles, so there is no chain to wait
// on: a CTA only waits for tiles that started before it (tile numbers are handed out by an
// atomic ticket, so "earlier tile" == "already running or finished": no deadlock).
__device__ __forceinline__ void dc_scan_body(const ScanGeom& g, int16_t* __restrict__ dc,
                                             unsigned long long* __restrict__ agg,
                                             uint32_t* __restrict__ ticket)
{
    __shared__ Dc4 warp_val[kDcTile / 32];
    __shared__ uint32_t warp_flag[kDcTile / 32];
    __shared__ Dc4 s_red[kDcTile / 32];
    __shared__ uint32_t s_tile;
    // the ticket does not depend on huff_write (the counter was cleared three kernels ago and
    // completion is transitive, kernels.hpp): take it while the predecessor is still draining
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
#ifdef JB2_TIMELINE
    Control* const tl_ctrl = reinterpret_cast<Control*>(reinterpret_cast<char*>(ticket) - offsetof(Control, dc_ticket));
#endif
    JB2_TL(tl_ctrl, 10, true);
    grid_dependency_wait();
    grid_dependency_trigger();
    JB2_TL(tl_ctrl, 11, true);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t tile_mcu0 = tile * kDcTile;
    const uint32_t ri = (uint32_t)g.restart_interval;
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
    if (threadIdx.x < 2) {
        const int k = threadIdx.x * 2;
        st_relaxed64(agg + 2 * (size_t)tile + threadIdx.x,
                     kDcValid | ((unsigned long long)(tail.v[k + 1] & 0xffff) << 16) |
                         (unsigned long long)(tail.v[k] & 0xffff));
    }
    // tiles [first_tile, tile) carry into this one
    const uint32_t seg_start = tile_mcu0 / ri * ri;
    const uint32_t first_tile = seg_start / kDcTile;
    Dc4 carry;
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a) carry.v[a] = 0;
    for (uint32_t t = first_tile + threadIdx.x; t < tile; t += kDcTile) {
        unsigned long long w0, w1;
        while (((w0 = ld_relaxed64(agg + 2 * (size_t)t)) & kDcValid) == 0) __nanosleep(20);
        while (((w1 = ld_relaxed64(agg + 2 * (size_t)t + 1)) & kDcValid) == 0) __nanosleep(20);
        carry.v[0] += (int)(w0 & 0xffffu);
        carry.v[1] += (int)((w0 >> 16) & 0xffffu);
        carry.v[2] += (int)(w1 & 0xffffu);
        carry.v[3] += (int)((w1 >> 16) & 0xffffu);
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
    if (!valid) {
        JB2_TL(tl_ctrl, 12, false);
        return;
    }
    // the tile carry applies only while no restart boundary has been seen inside the tile
    const bool fresh = (mcu / ri * ri) >= tile_mcu0;
    Dc4 pred; // predictor before this MCU
#pragma unroll
    for (int a = 0; a < kMaxComp; ++a)
        pred.v[a] = inc.v[a] - tot.v[a] + (fresh ? 0 : carry.v[a]);
    int16_t* base = dc + (size_t)mcu * g.dus_per_mcu;
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
            base[s] = (int16_t)value;
        }
    }
    JB2_TL(tl_ctrl, 12, false);
}

__global__ void __launch_bounds__(kDcTile)
dc_scan(const ScanGeom g, int16_t* __restrict__ dc, unsigned long long* __restrict__ agg,
        uint32_t* __restrict__ ticket)
{
    dc_scan_body(g, dc, agg, ticket);
}

__global__ void __launch_bounds__(kDcTile)
dc_scan_batch(const BatchImage* __restrict__ images)
{
    const BatchImage& im = images[blockIdx.y];
    if (blockIdx.x >= im.ctas[kBkDc]) return;
    __shared__ ScanGeom g;
    {
        static_assert(sizeof(ScanGeom) % 4 == 0, "copied word by word");
        const uint32_t* src = reinterpret_cast<const uint32_t*>(&im.g);
        uint32_t* dst = reinterpret_cast<uint32_t*>(&g);
        for (uint32_t i = threadIdx.x; i < sizeof(ScanGeom) / 4; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    dc_scan_body(g, im.b.dc, im.b.dc_agg, &im.b.ctrl->dc_ticket);
}

} // namespace

uint32_t dc_scan_ctas(const ScanGeom& g) { return ((uint32_t)g.total_mcus + kDcTile - 1) / kDcTile; }

cudaError_t launch_dc_batch(const BatchLaunch& bl, cudaStream_t stream)
{
    static PerDeviceOnce configured;
    configured.run([] { prefer_max_shared(dc_scan_batch); });
    return launch_dependent(dc_scan_batch, dim3(bl.max_ctas[kBkDc], bl.count), kDcTile, 0, stream, bl.d_images);
}

cudaError_t launch_dc(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm)
{
    if (g.total_mcus == 0) return cudaSuccess;
    static PerDeviceOnce configured;
    configured.run([] { prefer_max_shared(dc_scan); });
    const uint32_t tiles = dc_scan_ctas(g);
    tm->mark(stream, kIdDcScan);
    return launch_dependent(dc_scan, tiles, kDcTile, 0, stream, g, b.dc, b.dc_agg, &b.ctrl->dc_ticket);
}

} // namespace jb2
