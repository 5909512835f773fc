// kernels.hpp — launch wrappers of the sm_100a decode stages.  One function per reference
// stage: destuff (a1), huffman sync+write (a7-a11), dc (a12), transpose (a13),
// idct (a14).  All launches go to the caller's stream; none synchronises.
#ifndef JB2_KERNELS_HPP_
#define JB2_KERNELS_HPP_

#include "common.hpp"

#include <cuda_runtime.h>

#include <atomic>
#include <mutex>

namespace jb2 {

#ifndef JB2_SYNC_WARPS
#define JB2_SYNC_WARPS 4
#endif
constexpr int kHuffThreads = JB2_SYNC_WARPS * 32; // decoding threads of a huff_sync CTA (one image)
constexpr int kDcTile = 256;  // MCUs per CTA in the DC scan

// Device pointers of one scan (all inside the caller's scratch buffer).
struct ScanBufs {
    const uint8_t* scan_bytes; // stuffed entropy-coded bytes (uploaded); 16-byte aligned
    uint32_t scan_lead;        // bytes at scan_bytes that precede the scan (batched upload of whole
                               // files: the scan starts wherever it starts; 0 .. 15)
    SegDesc* segs;             // segment directory, written by destuff
    ScanInfo* info;            // totals of the scan, written by destuff
    unsigned long long* destuff_desc; // look-back descriptors of destuff (2 words per tile) followed
    uint32_t* destuff_ticket;         // by its ticket: zero when destuff starts (uploaded zeros, then
    uint32_t destuff_words;           // re-zeroed by huff_sync); destuff_words 32-bit words in all
    const uint32_t* lut_pool;  // uploaded
    const HuffSlow* slow;      // uploaded
    Control* ctrl;             // zeroed by destuff
    uint32_t* chunk_state;     // zeroed by destuff: one mailbox per sync warp
    uint32_t num_chunks;
    unsigned long long* dc_agg; // zeroed by destuff: two flagged aggregate words per DC tile
    unsigned long long* write_agg; // zeroed by destuff: slot aggregate per huff_write CTA
    uint32_t clear_words;      // Control + mailboxes + DC flags + write aggregates, contiguous
    uint32_t* destuffed;       // big-endian words, segment k at sub_off * kSubWords
    int16_t* coef;             // zeroed by huff_sync (or by huff_write, fill_in_write); DU-major coded order, 64 zig-zag-order int16 per DU (slot 0 unused)
    uint32_t fill_in_write;    // huff_write clears the coefficient buffer stretch by stretch (else huff_sync clears it)
    size_t coef_bytes;         // multiple of 16
    int16_t* dc;               // zeroed like coef; one DC value per DU, coded order: differences
                               // after huff_write, absolute values after dc_scan
    size_t dc_bytes;           // multiple of 16
    SubEntry* entries;         // N
    uint32_t* qpos;            // N: first slot (segment relative) of every subsequence
    uint32_t* seg_idx;         // N
};

// ---- batched launches (jpeggpu_b200 batch API): one launch of each kernel for many images -----
// A 2-D grid: blockIdx.y = image, blockIdx.x = CTA of that image; CTAs past the image's own count
// leave at once.  The descriptor array lives in device memory (uploaded with the batch).
enum BatchKernel { kBkDestuff = 0, kBkSync, kBkWrite, kBkDc, kBkIdct, kBkCount };
struct BatchImage {
    ScanGeom g;
    ScanBufs b;
    uint8_t* plane[kMaxComp]; // output planes, by scan component
    int pitch[kMaxComp];
    uint32_t q[kMaxComp][32]; // packed quantisers (kernels_idct.cu pack_quantisers)
    uint32_t ctas[kBkCount];  // CTAs of every kernel for this image
    uint32_t pad[3];
};
struct BatchLaunch {
    const BatchImage* d_images;
    uint32_t count;
    uint32_t max_ctas[kBkCount];
    uint32_t lut_pool_words_max;    // largest LUT pool / compact LUT set of the batch: sizes the
    uint32_t lut_compact_words_max; // dynamic shared memory of the Huffman kernels
};
// Fills im.ctas and im.q (host side); `write_warps` = warps per huff_write CTA for this batch.
void batch_image_grid(BatchImage& im, const uint8_t (&q)[kMaxComp][64], int write_warps);
int batch_write_warps(uint32_t lut_compact_words_max);
cudaError_t launch_batch(const BatchLaunch& bl, int write_warps, cudaStream_t stream);
// per-stage pieces of the above (one translation unit per stage)
uint32_t destuff_tiles(const ScanGeom& g, const ScanBufs& b);
uint32_t huff_sync_ctas(uint32_t num_subs, bool batch);
uint32_t huff_write_ctas(uint32_t num_subs, int warps);
uint32_t dc_scan_ctas(const ScanGeom& g);
uint32_t idct_ctas(const ScanGeom& g);
void idct_pack_quantisers(const uint8_t (&q)[kMaxComp][64], uint32_t (&out)[kMaxComp][32]);
cudaError_t launch_destuff_batch(const BatchLaunch& bl, cudaStream_t stream);
cudaError_t launch_huffman_batch(const BatchLaunch& bl, int write_warps, cudaStream_t stream);
cudaError_t launch_dc_batch(const BatchLaunch& bl, cudaStream_t stream);
cudaError_t launch_idct_batch(const BatchLaunch& bl, cudaStream_t stream);

// Optional per-kernel timing (jpeggpu_b200_set_profiling): an event is recorded on the launching
// stream before every kernel and one after the last; off by default and absent from timed runs.
enum KernelId {
    kIdDestuff = 0, kIdSync, kIdWrite, kIdDcScan, kIdIdct, kIdTranspose, kIdIdctRgbi,
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
// Optional output-side fusion: IDCT straight to interleaved RGB, and the same
// conversion (reference util/util.h:33-107) on existing planes.
cudaError_t launch_idct_rgbi(const ScanGeom& g, const ScanBufs& b, const RgbOut& out,
                             cudaStream_t stream, StageTimer* tm);
cudaError_t launch_ycc_to_rgbi(const PlanarIn& in, const RgbOut& out, cudaStream_t stream);

// Runs `f` once per device (launch wrappers set their kernel attributes there).  Concurrent first
// calls - the host threads of jpeggpu_b200_decode_batch all start at once - wait until the
// attributes are set: a thread that merely saw "somebody has started" could launch a kernel
// with 210 KB of dynamic shared memory before the opt-in has happened.
struct PerDeviceOnce {
    std::atomic<unsigned long long> done{0};
    std::mutex lock;
    template <class F>
    void run(F&& f)
    {
        int dev = 0;
        cudaGetDevice(&dev);
        const unsigned long long bit = 1ull << (dev & 63);
        if (done.load(std::memory_order_acquire) & bit) return;
        std::lock_guard<std::mutex> guard(lock);
        if (done.load(std::memory_order_relaxed) & bit) return;
        f();
        done.fetch_or(bit, std::memory_order_release);
    }
};

// Every kernel of the library asks for the same shared-memory carveout: when consecutive
// kernels prefer different L1/shared splits the SMs are drained and reconfigured between them,
// which showed up as ~25 us of idle time in front of each Huffman kernel.
template <class Kernel>
inline void prefer_max_shared(Kernel* k, int max_dynamic_bytes = 0)
{
    cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (max_dynamic_bytes)
        cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, max_dynamic_bytes);
}

// Programmatic dependent launch (sm_90+): every kernel after the first of a scan is launched
// with the programmatic-stream-serialization attribute, so its CTAs may start while the previous
// kernel is still draining.  Device side, a kernel does what needs no input from its predecessor
// (tables into shared memory, zero-fill), then grid_dependency_wait() - which returns once the
// predecessor grid has completed and its writes are visible - and only THEN
// grid_dependency_trigger(), which lets its own successor start its prologue.  Triggering after
// the wait makes completion transitive: a prologue never runs before the grand-predecessor is
// done.  Every thread of every kernel executes the wait.
#if defined(__CUDACC__)
// Timeline stamps (diagnostic build only): slot `idx` keeps the LATEST stamp, or with `first` the
// EARLIEST one (stored inverted so that both are an atomicMax over a zeroed word).
#ifdef JB2_TIMELINE
__device__ __forceinline__ void tl_stamp(Control* c, int idx, bool first)
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    const uint32_t v = (uint32_t)t;
    atomicMax(&c->tl[idx], first ? ~v : v);
}
#define JB2_TL(c, idx, first)                               \
    do {                                                    \
        if (threadIdx.x == 0 && (c)) tl_stamp((c), (idx), (first)); \
    } while (0)
#else
#define JB2_TL(c, idx, first) \
    do {                      \
    } while (0)
#endif
__device__ __forceinline__ void grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_dependency_trigger()
{
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
#endif

template <class... KArgs, class... Args>
inline cudaError_t launch_dependent(void (*kernel)(KArgs...), dim3 grid, unsigned block,
                                    size_t smem, cudaStream_t stream, Args&&... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#define JB2_LAUNCH(expr)                          \
    do {                                          \
        const cudaError_t jb2_e_ = (expr);        \
        if (jb2_e_ != cudaSuccess) return jb2_e_; \
    } while (0)

// Kernel launches issued per scan by each stage (for bench.py's gpu_launches claim).
int count_launches(const ScanGeom& g, bool with_transpose);

} // namespace jb2
#endif
