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

constexpr int kIdctThreads = 128;
#ifndef JB2_IDCT_MIN_CTAS
#define JB2_IDCT_MIN_CTAS 6
#endif

__global__ void __launch_bounds__(kIdctThreads, JB2_IDCT_MIN_CTAS)
idct_kernel(const ScanGeom g, const int16_t* __restrict__ coef, const int16_t* __restrict__ dc,
            const OutPlanes out)
{
    __shared__ int s_q[kMaxComp][64];
    for (int i = threadIdx.x; i < kMaxComp * 64; i += kIdctThreads)
        s_q[i >> 6][i & 63] = out.q[i >> 6][i & 63];
    grid_dependency_wait();
    grid_dependency_trigger();
    __syncthreads();
    DuPos pos;
    if (!locate(g, blockIdx.x * kIdctThreads + threadIdx.x, pos)) return;
    const int a = pos.comp;
    const uint4* src = reinterpret_cast<const uint4*>(coef + (size_t)pos.du * 64);
    // the buffer holds every data unit in zig-zag order: un-zig-zag while loading (register
    // indices are compile-time constants after unrolling)
    int v[8][8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const uint4 w = src[r];
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int lo = (int)(short)(ww[k] & 0xffffu);
            const int hi = (int)(short)(ww[k] >> 16);
            const int n0 = natural_of(r * 8 + 2 * k), n1 = natural_of(r * 8 + 2 * k + 1);
            // src/idct.cu:178-180: the product is stored as int16
            v[n0 >> 3][n0 & 7] = (int)(short)(lo * s_q[a][n0]);
            v[n1 >> 3][n1 & 7] = (int)(short)(hi * s_q[a][n1]);
        }
    }
    v[0][0] = (int)(short)((int)dc[pos.du] * s_q[a][0]); // absolute DC lives in the compact array
#pragma unroll
    for (int x = 0; x < 8; ++x)
        idct8(v[0][x], v[1][x], v[2][x], v[3][x], v[4][x], v[5][x], v[6][x], v[7][x]);
#pragma unroll
    for (int y = 0; y < 8; ++y)
        idct8(v[y][0], v[y][1], v[y][2], v[y][3], v[y][4], v[y][5], v[y][6], v[y][7]);

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

} // namespace

cudaError_t launch_idct(const ScanGeom& g, const ScanBufs& b, const OutPlanes& out,
                        cudaStream_t stream, StageTimer* tm)
{
    const uint32_t total = g.du_prefix[g.num_comp];
    if (total == 0) return cudaSuccess;
    static PerDeviceOnce configured;
    configured.run([] { prefer_max_shared(idct_kernel); });
    tm->mark(stream, kIdIdct);
    return launch_dependent(idct_kernel, (total + kIdctThreads - 1) / kIdctThreads, kIdctThreads, 0,
                            stream, g, b.coef, b.dc, out);
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

int count_launches(const ScanGeom& g, bool with_transpose)
{
    int n = 0;
    if (g.scan_bytes) n += 1;
    if (g.num_subs) n += 2;
    if (g.total_mcus) n += 1;
    n += 1; // idct
    if (with_transpose) n += 1;
    return n;
}

} // namespace jb2
