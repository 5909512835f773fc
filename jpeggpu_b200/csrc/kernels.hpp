// kernels.hpp — launch wrappers of the sm_100a decode stages.  One function per reference
// stage: destuff (a1), huffman sync+write (a7-a11), dc (a12), transpose (a13),
// idct (a14).  All launches go to the caller's stream; none synchronises.
#ifndef JB2_KERNELS_HPP_
#define JB2_KERNELS_HPP_

#include "common.hpp"

#include <cuda_runtime.h>

namespace jb2 {

constexpr int kDestuffThreads = 256;
constexpr int kDestuffBytesPerThread = 16;
constexpr int kDestuffTile = kDestuffThreads * kDestuffBytesPerThread; // stuffed bytes per CTA
constexpr int kHuffThreads = 128;
constexpr int kRepairPasses = 3;
constexpr int kDcTile = 256; // MCUs per CTA in the DC scan

// Device pointers of one scan (all inside the caller's scratch buffer).
struct ScanBufs {
    const uint8_t* scan_bytes; // stuffed entropy-coded bytes (uploaded)
    const SegDesc* segs;       // uploaded
    const uint16_t* lut_pool;  // uploaded
    const HuffSlow* slow;      // uploaded
    Control* ctrl;             // zeroed
    uint32_t* destuffed;       // zeroed; big-endian words, segment k at sub_off * kSubWords
    int16_t* coef;             // zeroed; DU-major coded order, 64 natural-order int16 per DU
    unsigned long long* tile_agg; // destuff tile aggregates
    SubEntry* entries;         // N
    uint32_t* qpos;            // N + 1 exclusive prefix of entries[].n
    uint32_t* seg_idx;         // N
    int* dc_agg;               // 4 per DC tile
};

cudaError_t launch_destuff(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream);
cudaError_t launch_huffman(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream);
cudaError_t launch_dc(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream);
cudaError_t launch_idct(const ScanGeom& g, const ScanBufs& b, const OutPlanes& out,
                        cudaStream_t stream);
cudaError_t launch_transpose(const ScanGeom& g, const ScanBufs& b, const CoefPlanes& planes,
                             cudaStream_t stream);

// Kernel launches issued per scan by each stage (for bench.py's gpu_launches claim).
int count_launches(const ScanGeom& g, bool with_transpose);

} // namespace jb2
#endif
