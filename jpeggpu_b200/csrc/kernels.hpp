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

// Optional per-kernel timing (jpeggpu_b200_set_profiling): an event is recorded on the launching
// stream before every kernel and one after the last; off by default and absent from timed runs.
enum KernelId {
    kIdMemset = 0, kIdDestuffCount, kIdDestuffApply, kIdSpeculate, kIdRepair0, kIdRepair1,
    kIdRepair2, kIdFixup, kIdSubScan, kIdWrite, kIdDcSums, kIdDcApply, kIdIdct, kIdTranspose,
    kIdCount
};

struct StageTimer {
    static constexpr int kMaxMarks = 64;
    cudaEvent_t ev[kMaxMarks];
    int id[kMaxMarks];
    int n = 0;
    bool on = false;
    bool created = false;
    void mark(cudaStream_t stream, int kernel_id)
    {
        if (!on || n >= kMaxMarks) return;
        id[n] = kernel_id;
        cudaEventRecord(ev[n], stream);
        ++n;
    }
};

cudaError_t launch_destuff(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm);
cudaError_t launch_huffman(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm);
cudaError_t launch_dc(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm);
cudaError_t launch_idct(const ScanGeom& g, const ScanBufs& b, const OutPlanes& out,
                        cudaStream_t stream, StageTimer* tm);
cudaError_t launch_transpose(const ScanGeom& g, const ScanBufs& b, const CoefPlanes& planes,
                             cudaStream_t stream, StageTimer* tm);

// Kernel launches issued per scan by each stage (for bench.py's gpu_launches claim).
int count_launches(const ScanGeom& g, bool with_transpose);

} // namespace jb2
#endif
