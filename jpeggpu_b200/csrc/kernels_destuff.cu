// kernels_destuff.cu — byte unstuffing + restart-segment placement (reference stage a1,
// src/decode_destuff.cu:37-148 + the three CUB scans at :220-316).
//
// One launch, 16-byte vector loads, no per-byte int arrays: each CTA takes one 4 KiB tile of
// stuffed bytes.  The tile's prefix (data bytes and restart markers before it) comes from the
// host walk that already visits every 0xff to find the end of the scan (parser.cpp walk_scan,
// the reference's src/reader.cpp:443-489), so tiles are independent: classify, block scan,
// compaction through shared memory, coalesced stores.  The kernel also clears the control block
// and the sync mailboxes of the Huffman stage and writes the zero padding behind every segment.
// Output is the destuffed scan stored as big-endian 32-bit words (byte d lands at d ^ 3) so the
// Huffman kernels read MSB-first words directly; segment k starts at segs[k].sub_off*kSubBytes.
// Algorithmic bytes: Sb read + Sd written.
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
destuff_kernel(const uint8_t* __restrict__ src, uint32_t n, const TileDesc* __restrict__ tiles,
               const SegDesc* __restrict__ segs, int num_segments, uint8_t* __restrict__ dst,
               uint32_t* __restrict__ clear, uint32_t clear_words)
{
    __shared__ uint32_t warp_sums[kDestuffThreads / 32];
    __shared__ uint8_t stage[kDestuffTile];
    __shared__ uint16_t seg_delta[kDestuffTile];

    // first kernel of a scan (launched behind the upload copies): let huff_sync start its
    // prologue (tables, zero-fill of the coefficient buffer) right away
    grid_dependency_wait();
    grid_dependency_trigger();
    for (uint32_t i = blockIdx.x * kDestuffThreads + threadIdx.x; i < clear_words;
         i += gridDim.x * kDestuffThreads)
        clear[i] = 0u;

    const TileDesc tile = tiles[blockIdx.x];
    const uint32_t kept_base = tile.kept_before;
    const uint32_t seg_base = tile.segs_before;

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
    if ((tile_total >> 16) == 0u && (int)seg_base < num_segments) {
        // No restart marker inside the tile (every tile of a scan without DRI, most tiles
        // otherwise): one segment, one contiguous run of output bytes.  Whole big-endian words
        // are assembled from the staged bytes and stored with 4-byte stores; only the run's
        // unaligned head and tail go byte by byte.
        const SegDesc sd = segs[seg_base];
        const uint32_t r0 = kept_base - sd.kept_before;               // index inside the segment
        const uint32_t d0 = r0 + sd.sub_off * (uint32_t)kSubBytes;    // output position of stage[0]
        const uint32_t d_end = d0 + tile_kept;
        const uint32_t w_begin = (d0 + 3u) & ~3u, w_end = d_end & ~3u; // whole words [w_begin, w_end)
        if (w_begin < w_end) {
            uint32_t* words = reinterpret_cast<uint32_t*>(dst);
            for (uint32_t w = w_begin + 4u * threadIdx.x; w < w_end; w += 4u * kDestuffThreads) {
                const uint32_t o = w - d0;
                words[w >> 2] = ((uint32_t)stage[o] << 24) | ((uint32_t)stage[o + 1] << 16) |
                                ((uint32_t)stage[o + 2] << 8) | (uint32_t)stage[o + 3];
            }
        }
        // head [d0, min(w_begin, d_end)) and tail [max(w_end, head end), d_end): < 8 bytes
        const uint32_t head_end = w_begin < d_end ? w_begin : d_end;
        const uint32_t tail_begin = w_end > head_end ? w_end : head_end;
        const uint32_t edge = (head_end - d0) + (d_end - tail_begin);
        if (threadIdx.x < edge) {
            const uint32_t d = threadIdx.x < head_end - d0 ? d0 + threadIdx.x
                                                           : tail_begin + (threadIdx.x - (head_end - d0));
            dst[d ^ 3u] = stage[d - d0];
        }
        if (tile_kept != 0u && r0 + tile_kept == sd.nbytes) { // the segment ends in this tile: pad
            const uint32_t pad_end = (d_end - 1u + (uint32_t)kSubBytes) / (uint32_t)kSubBytes * (uint32_t)kSubBytes;
            for (uint32_t z = d_end + threadIdx.x; z < pad_end; z += kDestuffThreads) dst[z ^ 3u] = 0;
        }
        return;
    }
    for (uint32_t o = threadIdx.x; o < tile_kept; o += kDestuffThreads) {
        const uint32_t s = seg_base + seg_delta[o];
        if ((int)s >= num_segments) continue; // cannot happen: host and device rules agree
        const SegDesc sd = segs[s];
        // data byte number k of the scan lands at k - kept_before + sub_off * kSubBytes
        const uint32_t r = kept_base + o - sd.kept_before; // index inside the segment
        const uint32_t d = r + sd.sub_off * (uint32_t)kSubBytes;
        dst[d ^ 3u] = stage[o];
        if (r + 1u == sd.nbytes) { // last byte of the segment: pad to a whole subsequence
            const uint32_t pad_end = (d + (uint32_t)kSubBytes) / (uint32_t)kSubBytes * (uint32_t)kSubBytes;
            for (uint32_t z = d + 1u; z < pad_end; ++z) dst[z ^ 3u] = 0;
        }
    }
}

} // namespace

cudaError_t launch_destuff(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm)
{
    if (g.scan_bytes == 0) return cudaSuccess;
    static PerDeviceOnce configured;
    configured.run([] { prefer_max_shared(destuff_kernel); });
    const uint32_t tiles = (g.scan_bytes + kDestuffTile - 1) / kDestuffTile;
    // the control block, the mailboxes and the DC flags are contiguous (planner): one clear range
    const uint32_t clear_words = b.clear_words;
    tm->mark(stream, kIdDestuff);
    destuff_kernel<<<tiles, kDestuffThreads, 0, stream>>>(
        b.scan_bytes, g.scan_bytes, b.tiles, b.segs, g.num_segments,
        reinterpret_cast<uint8_t*>(b.destuffed), reinterpret_cast<uint32_t*>(b.ctrl), clear_words);
    return cudaGetLastError();
}

} // namespace jb2
