// kernels_idct.cu — fused dequantise + 8x8 fixed-point IDCT + level shift + clamp, reading the
// DU-major coded-order coefficient buffer and writing straight into the caller's planes (reference
// stages a13+a14 fused: src/decode_transpose.cu:41-132 mapping + src/idct.cu:44-223 arithmetic),
// plus the stand-alone transpose that produces the reference's planar-raster coefficient layout
// (the bit-exact parity artefact).
//
// idct_kernel: one thread per data unit, threads ordered by (component, block row, block
// column) so that a warp writes 32 horizontally adjacent blocks: every row store instruction
// covers 256 contiguous bytes.  The 64 coefficients live in registers for both passes; there is
// no shared-memory transpose.  All components of the scan in one launch.
// Arithmetic is the reference's exactly (int32 butterflies, int16 truncation after dequantise,
// after the column pass and after the row pass) except that the quantiser is read as unsigned
// (reference quirk Q3: `int8_t qval`, src/idct.cu:179).
// Algorithmic bytes: 2*C coefficient bytes read + P sample bytes written.
#include "kernels.hpp"

#include <cstring>

namespace jb2 {
namespace {

// zig-zag index -> raster index inside a data unit (T.81 Figure A.6), foldable at compile time
__host__ __device__ constexpr int natural_of(int z)
{
    constexpr int t[64] = {JB2_NATURAL_ORDER};
    return t[z];
}
struct ZigzagOf { // raster index -> zig-zag index
    uint8_t v[64];
};
constexpr ZigzagOf make_zigzag_of()
{
    ZigzagOf r{};
    for (int z = 0; z < 64; ++z) r.v[natural_of(z)] = (uint8_t)z;
    return r;
}
__constant__ ZigzagOf c_zigzag_of = make_zigzag_of();

__device__ __forceinline__ int unfixh(int x) { return (int)(short)((x + 0x8000) >> 16); }
__device__ __forceinline__ int unfixo(int x) { return (x + 0x1000) >> 13; }

// src/idct.cu:50-95
__device__ __forceinline__ void idct8(int& v0, int& v1, int& v2, int& v3, int& v4, int& v5,
                                      int& v6, int& v7)
{
    constexpr int cos_1_4 = 0x5a82, sin_1_8 = 0x30fc, cos_1_8 = 0x7642, osin_1_16 = 0x063e,
                  osin_5_16 = 0x1a9b, ocos_1_16 = 0x1f63, ocos_5_16 = 0x11c7;
    const int tmp10 = (v0 + v4) * cos_1_4;
    const int tmp11 = (v0 - v4) * cos_1_4;
    const int tmp12 = v2 * sin_1_8 - v6 * cos_1_8;
    const int tmp13 = v6 * sin_1_8 + v2 * cos_1_8;
    const int tmp20 = tmp10 + tmp13, tmp21 = tmp11 + tmp12;
    const int tmp22 = tmp11 - tmp12, tmp23 = tmp10 - tmp13;
    const int tmp30 = unfixo((v3 + v5) * cos_1_4);
    const int tmp31 = unfixo((v3 - v5) * cos_1_4);
    const int w1 = v1 << 2, w7 = v7 << 2;
    const int tmp40 = w1 + tmp30, tmp41 = w7 + tmp31;
    const int tmp42 = w1 - tmp30, tmp43 = w7 - tmp31;
    const int tmp50 = tmp40 * ocos_1_16 + tmp41 * osin_1_16;
    const int tmp51 = tmp40 * osin_1_16 - tmp41 * ocos_1_16;
    const int tmp52 = tmp42 * ocos_5_16 + tmp43 * osin_5_16;
    const int tmp53 = tmp42 * osin_5_16 - tmp43 * ocos_5_16;
    v0 = unfixh(tmp20 + tmp50);
    v1 = unfixh(tmp21 + tmp53);
    v2 = unfixh(tmp22 + tmp52);
    v3 = unfixh(tmp23 + tmp51);
    v4 = unfixh(tmp23 - tmp51);
    v5 = unfixh(tmp22 - tmp52);
    v6 = unfixh(tmp21 - tmp53);
    v7 = unfixh(tmp20 - tmp50);
}

// bytes {sat(p0), sat(p1), sat(p2), sat(p3)}, p0 lowest, each saturated to 0..255
__device__ __forceinline__ uint32_t pack_sat_u8(int p0, int p1, int p2, int p3)
{
    uint32_t hi, d;
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(hi) : "r"(p3), "r"(p2), "r"(0));
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(p1), "r"(p0), "r"(hi));
    return d;
}

// Kernel parameters: destinations without the byte quantisers (those travel as QPairs).
struct PlaneDst {
    uint8_t* ptr[kMaxComp];
    int pitch[kMaxComp];
};
struct RgbDst {
    uint8_t* ptr;
    int pitch, width, height;
    int cy, cb, cr;
};

// Position of a plane-order thread: scan component, block coordinates and coded DU index.
struct DuPos {
    int comp, bx, by;
    uint32_t du;
};

__device__ __forceinline__ bool locate(const ScanGeom& g, uint32_t t, DuPos& pos)
{
    if (t >= g.du_prefix[g.num_comp]) return false;
    int a = 0;
#pragma unroll
    for (int k = 1; k < kMaxComp; ++k)
        if (k < g.num_comp && t >= g.du_prefix[k]) a = k;
    const uint32_t l = t - g.du_prefix[a];
    const int bxn = g.blocks_x[a];
    pos.comp = a;
    pos.by = (int)(l / (uint32_t)bxn);
    pos.bx = (int)(l - (uint32_t)pos.by * (uint32_t)bxn);
    // T.81 A.2.3: MCU-major, then component, then row-major DUs inside the MCU
    // (src/decode_transpose.cu:66-126 is the forward form of this mapping)
    const int sx = g.ssx[a], sy = g.ssy[a];
    const int mx = pos.bx / sx, ix = pos.bx - mx * sx;
    const int my = pos.by / sy, iy = pos.by - my * sy;
    pos.du = ((uint32_t)my * (uint32_t)g.mcus_x + (uint32_t)mx) * (uint32_t)g.dus_per_mcu +
             (uint32_t)(g.du_off[a] + iy * sx + ix);
    return true;
}

// Quantisers as the kernels want them: one 32-bit word per PAIR of zig-zag-consecutive
// coefficients, bytes {q[2j], 0, 0, q[2j+1]}, so that one dp2a instruction per coefficient
// extracts the signed 16-bit coefficient from its pair word AND multiplies it with its unsigned
// 8-bit quantiser (dp2a.lo uses bytes 0,1 of the second operand, dp2a.hi bytes 2,3; the zero
// bytes silence the other half).  8 x 16-byte shared-memory loads per data unit instead of 64
// scalar ones.
struct QPairs {
    uint32_t w[kMaxComp][32];
};
QPairs pack_quantisers(const uint8_t (&q)[kMaxComp][64])
{
    QPairs r;
    for (int a = 0; a < kMaxComp; ++a)
        for (int j = 0; j < 32; ++j)
            r.w[a][j] = (uint32_t)q[a][natural_of(2 * j)] | ((uint32_t)q[a][natural_of(2 * j + 1)] << 24);
    return r;
}
__device__ __forceinline__ int dp2a_lo_su(uint32_t a, uint32_t b)
{
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(0));
    return d;
}
__device__ __forceinline__ int dp2a_hi_su(uint32_t a, uint32_t b)
{
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(0));
    return d;
}
__device__ __forceinline__ void stage_quantisers(uint32_t (*s_q)[32], const QPairs& q)
{
    for (int i = threadIdx.x; i < kMaxComp * 32; i += blockDim.x) s_q[i >> 5][i & 31] = q.w[i >> 5][i & 31];
}

// Loads data unit `du` (zig-zag order), dequantises with the packed table `q` (16-byte aligned
// shared memory) and runs both IDCT passes in registers: v[y][x] is the sample before the level
// shift.
#ifndef JB2_IDCT_STAGED
#define JB2_IDCT_STAGED 1
#endif
// Experiment: fetch every data unit with ONE cp.async.bulk of 128 bytes,
// completed on an mbarrier per warp buffer, into a tile with a 144-byte pitch (conflict-free
// without a swizzle) instead of eight 16-byte cp.async per lane.  Measured on B200, 12 MP,
// idct_kernel: 22.5 us with cp.async, 26.7 us with bulk copies (48 MP 67.3 / 79.3, 2 MP 12.3 / 14.2;
// profiles/r2_r_idct_bulk.txt): 285 768 bulk copies of 128 bytes cost more than they save - the
// copy engine wants larger pieces, and a warp's 32 data units are 128-byte pieces 768 bytes apart
// (plane-order threads over an MCU-interleaved buffer).  Same output.  Off by default.
#ifndef JB2_IDCT_BULK
#define JB2_IDCT_BULK 0
#endif
#ifndef JB2_IDCT_EARLY_FETCH
#define JB2_IDCT_EARLY_FETCH 1
#endif
// Warp-cooperative fetch of the warp's 32 data units.  Thread-per-data-unit loads touch 32
// different 128-byte lines per instruction; here eight lanes fetch the eight 16-byte chunks of
// ONE data unit, so an instruction covers four whole data units (4 lines), and the chunks go to
// the owning thread through a 4 KB shared-memory tile per warp.  Chunk c of the data unit of
// lane d sits at d*128 + ((c ^ (d & 7)) * 16): conflict-free for the stores (a quarter-warp
// writes the eight chunks of one unit) and for the loads (eight lanes read chunk r of eight
// consecutive units).  Every lane of the warp must call this (invalid lanes fetch nothing).
__device__ __forceinline__ void stage_units(const int16_t* __restrict__ coef, uint32_t du, bool valid,
                                            uint4* tile)
{
    const int lane = threadIdx.x & 31;
    const int c = lane & 7, sub = lane >> 3;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int d = i * 4 + sub;
        const uint32_t du_d = __shfl_sync(0xffffffffu, du, d);
        const bool ok = __shfl_sync(0xffffffffu, (int)valid, d) != 0;
        if (ok)
            tile[d * 8 + (c ^ (d & 7))] =
                __ldg(reinterpret_cast<const uint4*>(coef + (size_t)du_d * 64) + c);
    }
    __syncwarp();
}

template <bool PITCH144 = false> // layout of `tile`: 144-byte pitch (bulk copies) or XOR swizzle
__device__ __forceinline__ void load_idct(const int16_t* __restrict__ coef,
                                          const int16_t* __restrict__ dc, uint32_t du,
                                          const uint32_t* q, int (&v)[8][8], const uint4* tile)
{
    const int lane = threadIdx.x & 31;
    const uint4* src = reinterpret_cast<const uint4*>(coef + (size_t)du * 64);
    const uint4* qsrc = reinterpret_cast<const uint4*>(q);
    // the buffer holds every data unit in zig-zag order: un-zig-zag while loading (register
    // indices are compile-time constants after unrolling)
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const uint4 w = PITCH144 ? tile[lane * 9 + r] : JB2_IDCT_STAGED ? tile[lane * 8 + (r ^ (lane & 7))] : src[r];
        const uint4 qq = qsrc[r];
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
        const uint32_t qw[4] = {qq.x, qq.y, qq.z, qq.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int n0 = natural_of(r * 8 + 2 * k), n1 = natural_of(r * 8 + 2 * k + 1);
            // src/idct.cu:178-180: the product is stored as int16
            v[n0 >> 3][n0 & 7] = (int)(short)dp2a_lo_su(ww[k], qw[k]);
            v[n1 >> 3][n1 & 7] = (int)(short)dp2a_hi_su(ww[k], qw[k]);
        }
    }
    v[0][0] = (int)(short)((int)dc[du] * (int)(q[0] & 0xffu)); // absolute DC lives in the compact array
#pragma unroll
    for (int x = 0; x < 8; ++x)
        idct8(v[0][x], v[1][x], v[2][x], v[3][x], v[4][x], v[5][x], v[6][x], v[7][x]);
#pragma unroll
    for (int y = 0; y < 8; ++y)
        idct8(v[y][0], v[y][1], v[y][2], v[y][3], v[y][4], v[y][5], v[y][6], v[y][7]);
}

constexpr int kIdctThreads = 128;
#ifndef JB2_IDCT_UNITS
#define JB2_IDCT_UNITS 2
#endif
constexpr int kIdctUnits = JB2_IDCT_STAGED ? JB2_IDCT_UNITS : 1; // data units per thread

#ifndef JB2_IDCT_MIN_CTAS
#define JB2_IDCT_MIN_CTAS 5
#endif

// cp.async flavour of stage_units: the 16-byte chunks go from global to shared memory without
// passing through registers, so the fetch of the next data unit overlaps the IDCT of this one.
__device__ __forceinline__ void stage_units_async(const int16_t* __restrict__ coef, uint32_t du,
                                                  bool valid, uint4* tile)
{
    const int lane = threadIdx.x & 31;
    const int c = lane & 7, sub = lane >> 3;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int d = i * 4 + sub;
        const uint32_t du_d = __shfl_sync(0xffffffffu, du, d);
        const bool ok = __shfl_sync(0xffffffffu, (int)valid, d) != 0;
        if (ok) {
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(tile + d * 8 + (c ^ (d & 7)));
            const void* src = reinterpret_cast<const uint4*>(coef + (size_t)du_d * 64) + c;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
        }
    }
}
#if JB2_IDCT_BULK
// Bulk flavour: lane l copies its own data unit (128 contiguous bytes) to tile + l * 144.
__device__ __forceinline__ void stage_units_bulk(const int16_t* __restrict__ coef, uint32_t du, bool valid,
                                                 uint4* tile, unsigned long long* bar)
{
    const int lane = threadIdx.x & 31;
    const uint32_t bar_addr = (uint32_t)__cvta_generic_to_shared(bar);
    const uint32_t n = __popc(__ballot_sync(0xffffffffu, valid));
    if (lane == 0)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(n * 128u) : "memory");
    __syncwarp();
    if (valid) {
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(tile + lane * 9);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(coef + (size_t)du * 64), "r"(128u), "r"(bar_addr)
                     : "memory");
    }
}
__device__ __forceinline__ void bulk_wait(unsigned long long* bar, uint32_t parity)
{
    const uint32_t bar_addr = (uint32_t)__cvta_generic_to_shared(bar);
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "JB2_IDCT_WAIT:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra JB2_IDCT_DONE;\n"
                 "bra JB2_IDCT_WAIT;\n"
                 "JB2_IDCT_DONE:\n"
                 "}" ::"r"(bar_addr), "r"(parity)
                 : "memory");
}
#endif
__device__ __forceinline__ void async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void async_wait_all_but_one() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// Dequantise + IDCT + level shift + clamp of one data unit and its eight row stores.
__device__ __forceinline__ void store_unit(const ScanGeom& g, const int16_t* __restrict__ coef,
                                           const int16_t* __restrict__ dc,
                                           const PlaneDst& out, const DuPos& pos, const uint32_t* q,
                                           const uint4* tile)
{
    const int a = pos.comp;
    int v[8][8];
    load_idct<JB2_IDCT_BULK != 0>(coef, dc, pos.du, q, v, tile);

    const int x0 = pos.bx * 8, y0 = pos.by * 8;
    const int size_x = g.size_x[a], size_y = g.size_y[a];
    if (x0 >= size_x || y0 >= size_y) return; // MCU padding
    uint8_t* dst = out.ptr[a] + (size_t)y0 * out.pitch[a] + x0;
    const bool full_row = x0 + 8 <= size_x && ((reinterpret_cast<uintptr_t>(dst) | (uintptr_t)out.pitch[a]) & 7u) == 0;
#pragma unroll
    for (int y = 0; y < 8; ++y) {
        if (y0 + y >= size_y) break;
        int s[8];
#pragma unroll
        for (int x = 0; x < 8; ++x) s[x] = (int)(short)(v[y][x] + 128); // src/idct.cu:216
        uint8_t* row = dst + (size_t)y * out.pitch[a];
        if (full_row) {
            // clamp(s, 0, 255) and byte packing in one instruction per two samples
            uint2 pk;
            pk.x = pack_sat_u8(s[0], s[1], s[2], s[3]);
            pk.y = pack_sat_u8(s[4], s[5], s[6], s[7]);
            *reinterpret_cast<uint2*>(row) = pk;
        } else {
#pragma unroll
            for (int x = 0; x < 8; ++x)
                if (x0 + x < size_x) row[x] = (uint8_t)min(max(s[x], 0), 255);
        }
    }
}

// s_q: the packed quantisers in shared memory (staged by the caller, before the barrier below)
__device__ __forceinline__ void idct_body(const ScanGeom& g, const int16_t* __restrict__ coef,
                                          const int16_t* __restrict__ dc, const PlaneDst& out,
                                          uint32_t (*s_q)[32], const uint32_t block)
{
#if JB2_IDCT_STAGED
    // Software pipeline over kIdctUnits data units per thread: while the IDCT of one unit runs,
    // the cp.async fetch of the next one is in flight (double-buffered tile per warp).
    constexpr int kTileVecs = JB2_IDCT_BULK ? 32 * 9 : 256;
    __shared__ uint4 s_tile[kIdctThreads / 32][2][kTileVecs];
    uint4* const tiles = s_tile[threadIdx.x >> 5][0];
    const uint32_t first = block * (uint32_t)(kIdctThreads * kIdctUnits) + threadIdx.x;
    DuPos pos, nxt;
    pos.comp = nxt.comp = 0;
    pos.du = nxt.du = 0;
    bool located = locate(g, first, pos);
#if !JB2_IDCT_BULK && JB2_IDCT_EARLY_FETCH
    // The coefficients come from huff_write, the grand-predecessor: it is complete when this CTA
    // starts (dc_scan triggers its dependents only behind its own dependency wait), so the fetch of
    // the first data unit can run while dc_scan - which only produces the DC values - is at work.
    stage_units_async(coef, pos.du, located, tiles);
    async_commit();
#endif
    grid_dependency_wait();
    grid_dependency_trigger();
    __syncthreads();
#if JB2_IDCT_BULK
    __shared__ __align__(8) unsigned long long s_bar[kIdctThreads / 32][2];
    unsigned long long* const bars = s_bar[threadIdx.x >> 5];
    if ((threadIdx.x & 31) == 0) {
        const uint32_t b0 = (uint32_t)__cvta_generic_to_shared(bars);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b0) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b0 + 8u) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async;" ::: "memory");
    }
    __syncwarp();
#endif
#if JB2_IDCT_BULK
    stage_units_bulk(coef, pos.du, located, tiles, bars);
#elif !JB2_IDCT_EARLY_FETCH
    stage_units_async(coef, pos.du, located, tiles);
    async_commit();
#endif
#pragma unroll 1
    for (int it = 0; it < kIdctUnits; ++it) {
        const uint4* tile = tiles + (it & 1) * kTileVecs;
        bool located_next = false;
#if JB2_IDCT_BULK
        if (it + 1 < kIdctUnits) {
            located_next = locate(g, first + (uint32_t)(it + 1) * kIdctThreads, nxt);
            stage_units_bulk(coef, nxt.du, located_next, tiles + ((it + 1) & 1) * kTileVecs, bars + ((it + 1) & 1));
        }
        bulk_wait(bars + (it & 1), (uint32_t)(it >> 1) & 1u);
#else
        if (it + 1 < kIdctUnits) {
            located_next = locate(g, first + (uint32_t)(it + 1) * kIdctThreads, nxt);
            stage_units_async(coef, nxt.du, located_next, tiles + ((it + 1) & 1) * kTileVecs);
        }
        async_commit();
        async_wait_all_but_one();
        __syncwarp();
#endif
        if (located) store_unit(g, coef, dc, out, pos, s_q[pos.comp], tile);
        __syncwarp(); // the tile is refilled two iterations from now
        pos = nxt;
        located = located_next;
    }
#else
    grid_dependency_wait();
    grid_dependency_trigger();
    __syncthreads();
    DuPos pos;
    if (!locate(g, block * kIdctThreads + threadIdx.x, pos)) return;
    store_unit(g, coef, dc, out, pos, s_q[pos.comp], nullptr);
#endif
}

__global__ void __launch_bounds__(kIdctThreads, JB2_IDCT_MIN_CTAS)
idct_kernel(const ScanGeom g, const int16_t* __restrict__ coef, const int16_t* __restrict__ dc,
            const PlaneDst out, const QPairs q)
{
    __shared__ __align__(16) uint32_t s_q[kMaxComp][32];
    stage_quantisers(s_q, q);
    idct_body(g, coef, dc, out, s_q, blockIdx.x);
}

__global__ void __launch_bounds__(kIdctThreads, JB2_IDCT_MIN_CTAS)
idct_batch_kernel(const BatchImage* __restrict__ images)
{
    const BatchImage& im = images[blockIdx.y];
    if (blockIdx.x >= im.ctas[kBkIdct]) return;
    __shared__ __align__(16) uint32_t s_q[kMaxComp][32];
    __shared__ ScanGeom g;
    __shared__ PlaneDst out;
    {
        static_assert(sizeof(ScanGeom) % 4 == 0, "copied word by word");
        const uint32_t* src = reinterpret_cast<const uint32_t*>(&im.g);
        uint32_t* dst = reinterpret_cast<uint32_t*>(&g);
        for (uint32_t i = threadIdx.x; i < sizeof(ScanGeom) / 4; i += blockDim.x) dst[i] = __ldg(src + i);
        for (int i = threadIdx.x; i < kMaxComp * 32; i += blockDim.x) s_q[i >> 5][i & 31] = __ldg(&im.q[i >> 5][i & 31]);
        if (threadIdx.x < kMaxComp) {
            out.ptr[threadIdx.x] = im.plane[threadIdx.x];
            out.pitch[threadIdx.x] = im.pitch[threadIdx.x];
        }
    }
    __syncthreads(); // idct_body reads the geometry before its own barrier (early fetch)
    idct_body(g, im.b.coef, im.b.dc, out, s_q, blockIdx.x);
}

// Stand-alone transpose: coded order (zig-zag inside a DU) -> planar raster int16 (row length =
// plane stride).  One thread gathers one 8-coefficient raster row, eight consecutive threads one DU.
__global__ void __launch_bounds__(256)
transpose_kernel(const ScanGeom g, const int16_t* __restrict__ coef,
                 const int16_t* __restrict__ dc, const CoefPlanes planes)
{
    grid_dependency_wait();
    grid_dependency_trigger();
    const uint32_t t = blockIdx.x * 256 + threadIdx.x;
    const uint32_t r = t & 7u;
    DuPos pos;
    if (!locate(g, t >> 3, pos)) return;
    const int16_t* du = coef + (size_t)pos.du * 64;
    uint32_t px[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t lo = (uint16_t)du[c_zigzag_of.v[r * 8 + 2 * k]];
        const uint32_t hi = (uint16_t)du[c_zigzag_of.v[r * 8 + 2 * k + 1]];
        px[k] = lo | (hi << 16);
    }
    if (r == 0) px[0] = (px[0] & 0xffff0000u) | (uint16_t)dc[pos.du];
    int16_t* dst = planes.ptr[pos.comp] +
                   (size_t)(pos.by * 8 + (int)r) * planes.stride[pos.comp] + pos.bx * 8;
    *reinterpret_cast<uint4*>(dst) = make_uint4(px[0], px[1], px[2], px[3]);
}

// ---- optional output-side fusion: chroma replication + YCbCr -> interleaved RGB ----------------
// (the step the reference's callers run on the CPU, util/util.h:33-107.)
// The arithmetic is the reference's float expression, operation for operation, with explicit
// round-to-nearest multiplies and adds so that nvcc cannot contract them into FMAs (a CPU build
// of util.h without FMA contraction is the parity target): r = y + 1.402 (cr-128),
// g = (y - .344136 (cb-128)) - .714136 (cr-128), b = y + 1.772 (cb-128), roundf, clamp 0..255.
struct ChromaTerms {
    float rr, gb, gr, bb;
};
// None of the conversions below touches the (eighth-rate) I2F/F2I pipe.  Integers 0..255 become
// floats by planting them in the mantissa of 2^23; round-half-away + clamp of the result is
// floor(x + 0.5) taken by a round-down multiply into the denormal range, where the bit pattern IS
// the integer (negative inputs give a negative pattern), and the saturating byte pack clamps.
// floor(fl(x + 0.5)) equals clamp(roundf(x)) for every value the matrix can produce from 8-bit
// inputs; tests/test_gpu_rgbi.py::test_convert_rgbi_every_triple checks all 2^24 triples.
__device__ __forceinline__ float u8_to_float(int v, float bias)
{
    return __fsub_rn(__int_as_float(0x4b000000 | v), bias); // (2^23 + v) - bias, exact
}
__device__ __forceinline__ ChromaTerms chroma_terms(int cb, int cr)
{
    const float fb = u8_to_float(cb, 8388736.f), fr = u8_to_float(cr, 8388736.f); // c - 128
    ChromaTerms t;
    t.rr = __fmul_rn(1.402f, fr);
    t.gb = __fmul_rn(.344136f, fb);
    t.gr = __fmul_rn(.714136f, fr);
    t.bb = __fmul_rn(1.772f, fb);
    return t;
}
__device__ __forceinline__ int round_away(float x)
{
    return __float_as_int(__fmul_rd(__fadd_rn(x, 0.5f), __int_as_float(1))); // * 2^-149, toward -inf
}
// cy must already be clamped to 0..255; r, g, b come back unclamped (negative: any negative int)
__device__ __forceinline__ void ycc_pixel(int cy, const ChromaTerms& t, int& r, int& g, int& b)
{
    const float fy = u8_to_float(cy, 8388608.f);
    r = round_away(__fadd_rn(fy, t.rr));
    g = round_away(__fsub_rn(__fsub_rn(fy, t.gb), t.gr));
    b = round_away(__fadd_rn(fy, t.bb));
}

// Stores pixels [0, n) of eight RGB triples at `row`; vec: row is 8-byte aligned and n == 8.
__device__ __forceinline__ void store_rgb8(uint8_t* row, const int (&r)[8], const int (&g)[8],
                                           const int (&b)[8], bool vec, int n)
{
    if (vec) {
        uint2* d = reinterpret_cast<uint2*>(row);
        d[0] = make_uint2(pack_sat_u8(r[0], g[0], b[0], r[1]), pack_sat_u8(g[1], b[1], r[2], g[2]));
        d[1] = make_uint2(pack_sat_u8(b[2], r[3], g[3], b[3]), pack_sat_u8(r[4], g[4], b[4], r[5]));
        d[2] = make_uint2(pack_sat_u8(g[5], b[5], r[6], g[6]), pack_sat_u8(b[6], r[7], g[7], b[7]));
    } else {
#pragma unroll
        for (int x = 0; x < 8; ++x)
            if (x < n) {
                row[3 * x + 0] = (uint8_t)min(max(r[x], 0), 255);
                row[3 * x + 1] = (uint8_t)min(max(g[x], 0), 255);
                row[3 * x + 2] = (uint8_t)min(max(b[x], 0), 255);
            }
    }
}

// idct_rgbi_kernel: fused dequantise + IDCT + chroma replication + colour conversion for an image
// that is ONE interleaved scan of Y (ssx x ssy data units per MCU) and 1x1 Cb, Cr - or a single
// grey component.  A CTA owns `mcus` horizontally adjacent MCUs.  Thread order: the Y data units
// first, by (row inside the MCU, MCU, column inside the MCU) so that a warp holds horizontally
// adjacent blocks and its three 8-byte stores per row cover 768 contiguous bytes; then `mcus` Cb
// threads and `mcus` Cr threads.  Every thread runs the IDCT of its data unit; the chroma threads
// leave their clamped samples in shared memory (72-byte stride per data unit: conflict-free 8-byte
// row reads), the Y threads convert and store.  The planes are never written.
// Algorithmic bytes: 2*C coefficient bytes read + 3*W*H bytes written.
#ifndef JB2_RGB_MCUS
#define JB2_RGB_MCUS 32
#endif
constexpr int kRgbMcus = JB2_RGB_MCUS; // MCUs per CTA of a colour image
constexpr int kRgbGreyMcus = 128; // data units per CTA of a grey image
constexpr int kRgbMaxThreads = kRgbMcus * kMaxDuPerMcu;

#ifndef JB2_RGB_MIN_CTAS
#define JB2_RGB_MIN_CTAS 3
#endif
// SSX, SSY: Y data units per MCU in x / y for the common layouts 1x1 (4:4:4, grey), 2x1 (4:2:2)
// and 2x2 (4:2:0) - the CTA size and the register budget follow from them; 0, 0: any layout
// (runtime g.ssx / g.ssy, CTA of up to kRgbMaxThreads).
template <int SSX, int SSY>
__global__ void __launch_bounds__(SSX ? kRgbMcus * (SSX * SSY + 2) : kRgbMaxThreads,
                                  SSX ? JB2_RGB_MIN_CTAS : 1)
idct_rgbi_kernel(const ScanGeom g, const int16_t* __restrict__ coef, const int16_t* __restrict__ dc,
                 const RgbDst out, const QPairs q)
{
    __shared__ __align__(16) uint32_t s_q[kMaxComp][32];
    __shared__ unsigned long long s_c[2][kRgbMcus][9];
    stage_quantisers(s_q, q);
    grid_dependency_wait();
    grid_dependency_trigger();
    __syncthreads();

    const bool grey = out.cb < 0;
    const int mcus = grey ? kRgbGreyMcus : kRgbMcus;
    const int ssx = SSX ? SSX : g.ssx[out.cy], ssy = SSY ? SSY : g.ssy[out.cy];
    const int groups_x = (g.mcus_x + mcus - 1) / mcus;
    const int my = blockIdx.x / groups_x;
    const int mx0 = (blockIdx.x - my * groups_x) * mcus;
    const int t = threadIdx.x;
    const int y_threads = mcus * ssx * ssy;
    const bool is_y = t < y_threads;
    int a, m, ix = 0, iy = 0, k = 0;
    if (is_y) {
        const int per_row = mcus * ssx;
        iy = t / per_row;
        const int rem = t - iy * per_row;
        m = rem / ssx;
        ix = rem - m * ssx;
        a = out.cy;
    } else {
        k = (t - y_threads) / mcus;
        m = (t - y_threads) - k * mcus;
        a = k ? out.cr : out.cb;
    }
    const int mx = mx0 + m;
    const bool valid = mx < g.mcus_x;
    int v[8][8];
    const uint32_t du = valid ? ((uint32_t)my * (uint32_t)g.mcus_x + (uint32_t)mx) * (uint32_t)g.dus_per_mcu +
                                    (uint32_t)(g.du_off[a] + iy * ssx + ix)
                              : 0u;
#if JB2_IDCT_STAGED
    __shared__ uint4 s_tile[kRgbMaxThreads / 32][256];
    uint4* tile = s_tile[threadIdx.x >> 5];
    stage_units(coef, du, valid, tile);
#else
    const uint4* tile = nullptr;
#endif
    if (valid) {
        load_idct(coef, dc, du, s_q[a], v, tile);
        if (!is_y) {
#pragma unroll
            for (int y = 0; y < 8; ++y) {
                int s[8];
#pragma unroll
                for (int x = 0; x < 8; ++x) s[x] = (int)(short)(v[y][x] + 128);
                s_c[k][m][y] = (unsigned long long)pack_sat_u8(s[0], s[1], s[2], s[3]) |
                               ((unsigned long long)pack_sat_u8(s[4], s[5], s[6], s[7]) << 32);
            }
        }
    }
    if (!grey) __syncthreads();
    if (!is_y || !valid) return;

    const int x0 = (mx * ssx + ix) * 8, y0 = (my * ssy + iy) * 8;
    if (x0 >= out.width || y0 >= out.height) return; // MCU padding
    uint8_t* dst = out.ptr + (size_t)y0 * out.pitch + (size_t)x0 * 3;
    const int n = min(8, out.width - x0);
    const bool vec = n == 8 && ((reinterpret_cast<uintptr_t>(out.ptr) | (uintptr_t)out.pitch) & 7u) == 0;
#pragma unroll
    for (int y = 0; y < 8; ++y) {
        if (y0 + y >= out.height) break;
        int s[8], r[8], gg[8], b[8];
#pragma unroll
        for (int x = 0; x < 8; ++x) s[x] = __vimin_s32_relu((int)(short)(v[y][x] + 128), 255);
        if (grey) {
#pragma unroll
            for (int x = 0; x < 8; ++x) r[x] = gg[x] = b[x] = s[x];
        } else {
            const int yc = (iy * 8 + y) / ssy;
            const unsigned long long cbrow = s_c[0][m][yc], crrow = s_c[1][m][yc];
            if (SSX == 1 || SSX == 2) {
                constexpr int S = SSX ? SSX : 1;
                // chroma column of pixel x: ix * (8 / SSX) + x / SSX
                const int sh = ix * (64 / S);
                const uint32_t cbw = (uint32_t)(cbrow >> sh), crw = (uint32_t)(crrow >> sh);
                const uint32_t cbw2 = (uint32_t)(cbrow >> 32), crw2 = (uint32_t)(crrow >> 32);
#pragma unroll
                for (int j = 0; j < 8 / S; ++j) {
                    const uint32_t wb = (S == 1 && j >= 4) ? cbw2 : cbw;
                    const uint32_t wr = (S == 1 && j >= 4) ? crw2 : crw;
                    const ChromaTerms ct = chroma_terms((wb >> (8 * (j & 3))) & 0xff, (wr >> (8 * (j & 3))) & 0xff);
#pragma unroll
                    for (int e = 0; e < S; ++e)
                        ycc_pixel(s[j * S + e], ct, r[j * S + e], gg[j * S + e], b[j * S + e]);
                }
            } else {
#pragma unroll
                for (int x = 0; x < 8; ++x) {
                    const int xc = (ix * 8 + x) / ssx;
                    const ChromaTerms ct = chroma_terms((int)(cbrow >> (8 * xc)) & 0xff, (int)(crrow >> (8 * xc)) & 0xff);
                    ycc_pixel(s[x], ct, r[x], gg[x], b[x]);
                }
            }
        }
        store_rgb8(dst + (size_t)y * out.pitch, r, gg, b, vec, n);
    }
}

// ycc_to_rgbi_kernel: the same conversion on planes that already exist (any scan structure):
// one thread per group of eight horizontally adjacent output pixels.  LG = log2 of the luma
// samples per chroma sample in x when that is 1, 2 or 4 (vector loads of the eight luma bytes and
// of the 8 >> LG chroma bytes when the planes are aligned, chroma terms shared between the pixels
// of one chroma sample); LG < 0: any factor, byte loads.
// Algorithmic bytes: P plane bytes read + 3*W*H written.
template <int LG>
__global__ void __launch_bounds__(256)
ycc_to_rgbi_kernel(const PlanarIn in, const RgbDst out)
{
    const int groups_x = (out.width + 7) >> 3;
    const uint32_t t = blockIdx.x * 256u + threadIdx.x;
    const int y = (int)(t / (uint32_t)groups_x);
    if (y >= out.height) return;
    const int x0 = (int)(t - (uint32_t)y * (uint32_t)groups_x) * 8;
    const int n = min(8, out.width - x0);
    const uint8_t* py = in.ptr[0] + (size_t)y * in.pitch[0] + x0;
    const bool grey = in.num_comp == 1;
    int s[8], r[8], gg[8], b[8];
    if (n == 8 && in.aligned) {
        const uint2 w = *reinterpret_cast<const uint2*>(py);
#pragma unroll
        for (int x = 0; x < 8; ++x) s[x] = ((x < 4 ? w.x : w.y) >> (8 * (x & 3))) & 0xff;
    } else {
#pragma unroll
        for (int x = 0; x < 8; ++x) s[x] = x < n ? py[x] : 0;
    }
    if (grey) {
#pragma unroll
        for (int x = 0; x < 8; ++x) r[x] = gg[x] = b[x] = s[x];
    } else {
        const uint8_t* pb = in.ptr[1] + (size_t)(y / in.ssy) * in.pitch[1];
        const uint8_t* pr = in.ptr[2] + (size_t)(y / in.ssy) * in.pitch[2];
        if (LG >= 0) {
            constexpr int L = LG >= 0 ? LG : 0;
            constexpr int kC = 8 >> L; // chroma samples under eight luma pixels
            const int c0 = x0 >> L;
            unsigned long long cbrow = 0, crrow = 0;
            if (n == 8 && in.aligned) {
                if (kC == 8) {
                    cbrow = *reinterpret_cast<const unsigned long long*>(pb + c0);
                    crrow = *reinterpret_cast<const unsigned long long*>(pr + c0);
                } else if (kC == 4) {
                    cbrow = *reinterpret_cast<const uint32_t*>(pb + c0);
                    crrow = *reinterpret_cast<const uint32_t*>(pr + c0);
                } else {
                    cbrow = *reinterpret_cast<const uint16_t*>(pb + c0);
                    crrow = *reinterpret_cast<const uint16_t*>(pr + c0);
                }
            } else {
#pragma unroll
                for (int j = 0; j < kC; ++j)
                    if ((j << L) < n) {
                        cbrow |= (unsigned long long)pb[c0 + j] << (8 * j);
                        crrow |= (unsigned long long)pr[c0 + j] << (8 * j);
                    }
            }
#pragma unroll
            for (int j = 0; j < kC; ++j) {
                const ChromaTerms ct = chroma_terms((int)(cbrow >> (8 * j)) & 0xff, (int)(crrow >> (8 * j)) & 0xff);
#pragma unroll
                for (int e = 0; e < (1 << L); ++e)
                    ycc_pixel(s[(j << L) + e], ct, r[(j << L) + e], gg[(j << L) + e], b[(j << L) + e]);
            }
        } else {
#pragma unroll
            for (int x = 0; x < 8; ++x) {
                r[x] = gg[x] = b[x] = 0;
                if (x < n) {
                    const int xc = (x0 + x) / in.ssx;
                    ycc_pixel(s[x], chroma_terms(pb[xc], pr[xc]), r[x], gg[x], b[x]);
                }
            }
        }
    }
    const bool vec = n == 8 && ((reinterpret_cast<uintptr_t>(out.ptr) | (uintptr_t)out.pitch) & 7u) == 0;
    store_rgb8(out.ptr + (size_t)y * out.pitch + (size_t)x0 * 3, r, gg, b, vec, n);
}

} // namespace

uint32_t idct_ctas(const ScanGeom& g)
{
    const uint32_t per_cta = (uint32_t)(kIdctThreads * kIdctUnits);
    return (g.du_prefix[g.num_comp] + per_cta - 1) / per_cta;
}

void idct_pack_quantisers(const uint8_t (&q)[kMaxComp][64], uint32_t (&out)[kMaxComp][32])
{
    const QPairs p = pack_quantisers(q);
    memcpy(out, p.w, sizeof(p.w));
}

cudaError_t launch_idct_batch(const BatchLaunch& bl, cudaStream_t stream)
{
    static PerDeviceOnce configured;
    configured.run([] { prefer_max_shared(idct_batch_kernel); });
    return launch_dependent(idct_batch_kernel, dim3(bl.max_ctas[kBkIdct], bl.count), kIdctThreads, 0, stream,
                            bl.d_images);
}

// Host side of a batched launch: the CTAs every kernel needs for one image, its packed quantisers.
void batch_image_grid(BatchImage& im, const uint8_t (&q)[kMaxComp][64], int write_warps)
{
    im.ctas[kBkDestuff] = destuff_tiles(im.g, im.b);
    im.ctas[kBkSync] = huff_sync_ctas(im.g.num_subs, true);
    im.ctas[kBkWrite] = huff_write_ctas(im.g.num_subs, write_warps);
    im.ctas[kBkDc] = dc_scan_ctas(im.g);
    im.ctas[kBkIdct] = idct_ctas(im.g);
    idct_pack_quantisers(q, im.q);
}

// destuff -> huff_sync -> huff_write -> dc_scan -> idct for every image of the batch: five launches.
cudaError_t launch_batch(const BatchLaunch& bl, int write_warps, cudaStream_t stream)
{
    if (bl.count == 0) return cudaSuccess;
    JB2_LAUNCH(launch_destuff_batch(bl, stream));
    JB2_LAUNCH(launch_huffman_batch(bl, write_warps, stream));
    JB2_LAUNCH(launch_dc_batch(bl, stream));
    return launch_idct_batch(bl, stream);
}

cudaError_t launch_idct(const ScanGeom& g, const ScanBufs& b, const OutPlanes& out,
                        cudaStream_t stream, StageTimer* tm)
{
    const uint32_t total = g.du_prefix[g.num_comp];
    if (total == 0) return cudaSuccess;
    static PerDeviceOnce configured;
    configured.run([] { prefer_max_shared(idct_kernel); });
    tm->mark(stream, kIdIdct);
    PlaneDst dst;
    for (int a = 0; a < kMaxComp; ++a) {
        dst.ptr[a] = out.ptr[a];
        dst.pitch[a] = out.pitch[a];
    }
    return launch_dependent(idct_kernel, idct_ctas(g), kIdctThreads, 0,
                            stream, g, b.coef, b.dc, dst, pack_quantisers(out.q));
}

cudaError_t launch_transpose(const ScanGeom& g, const ScanBufs& b, const CoefPlanes& planes,
                             cudaStream_t stream, StageTimer* tm)
{
    const uint32_t total = g.du_prefix[g.num_comp] * 8u;
    if (total == 0) return cudaSuccess;
    static PerDeviceOnce configured;
    configured.run([] { prefer_max_shared(transpose_kernel); });
    tm->mark(stream, kIdTranspose);
    return launch_dependent(transpose_kernel, (total + 255) / 256, 256, 0, stream, g, b.coef, b.dc,
                            planes);
}

static RgbDst rgb_dst(const RgbOut& out)
{
    RgbDst d;
    d.ptr = out.ptr;
    d.pitch = out.pitch;
    d.width = out.width;
    d.height = out.height;
    d.cy = out.cy;
    d.cb = out.cb;
    d.cr = out.cr;
    return d;
}

cudaError_t launch_idct_rgbi(const ScanGeom& g, const ScanBufs& b, const RgbOut& out,
                             cudaStream_t stream, StageTimer* tm)
{
    if (g.total_mcus == 0) return cudaSuccess;
    const bool grey = out.cb < 0;
    const int mcus = grey ? kRgbGreyMcus : kRgbMcus;
    const int ssx = g.ssx[out.cy], ssy = g.ssy[out.cy];
    const unsigned threads = grey ? kRgbGreyMcus : kRgbMcus * (ssx * ssy + 2);
    const unsigned grid = (unsigned)((g.mcus_x + mcus - 1) / mcus) * (unsigned)g.mcus_y;
    static PerDeviceOnce configured;
    configured.run([] {
        prefer_max_shared(idct_rgbi_kernel<0, 0>);
        prefer_max_shared(idct_rgbi_kernel<1, 1>);
        prefer_max_shared(idct_rgbi_kernel<2, 1>);
        prefer_max_shared(idct_rgbi_kernel<2, 2>);
    });
    tm->mark(stream, kIdIdctRgbi);
    auto* kernel = grey ? idct_rgbi_kernel<0, 0>
                   : ssx == 1 && ssy == 1 ? idct_rgbi_kernel<1, 1>
                   : ssx == 2 && ssy == 1 ? idct_rgbi_kernel<2, 1>
                   : ssx == 2 && ssy == 2 ? idct_rgbi_kernel<2, 2> : idct_rgbi_kernel<0, 0>;
    return launch_dependent(kernel, grid, threads, 0, stream, g, b.coef, b.dc, rgb_dst(out),
                            pack_quantisers(out.q));
}

cudaError_t launch_ycc_to_rgbi(const PlanarIn& in, const RgbOut& out, cudaStream_t stream)
{
    const unsigned long long groups = (unsigned long long)((out.width + 7) / 8) * (unsigned long long)out.height;
    if (groups == 0) return cudaSuccess;
    const unsigned grid = (unsigned)((groups + 255) / 256);
    PlanarIn p = in;
    const RgbDst dst = rgb_dst(out);
    uintptr_t bits = 0;
    for (int c = 0; c < in.num_comp; ++c) bits |= reinterpret_cast<uintptr_t>(in.ptr[c]) | (uintptr_t)in.pitch[c];
    p.aligned = (bits & 7u) == 0;
    switch (in.ssx) {
    case 1: ycc_to_rgbi_kernel<0><<<grid, 256, 0, stream>>>(p, dst); break;
    case 2: ycc_to_rgbi_kernel<1><<<grid, 256, 0, stream>>>(p, dst); break;
    case 4: ycc_to_rgbi_kernel<2><<<grid, 256, 0, stream>>>(p, dst); break;
    default: ycc_to_rgbi_kernel<-1><<<grid, 256, 0, stream>>>(p, dst); break;
    }
    return cudaGetLastError();
}

int count_launches(const ScanGeom& g, bool with_transpose)
{
    int n = 0;
    if (g.scan_avail) n += 1;
    if (g.num_subs) n += 2;
    if (g.total_mcus) n += 1;
    n += 1; // idct
    if (with_transpose) n += 1;
    return n;
}

} // namespace jb2
