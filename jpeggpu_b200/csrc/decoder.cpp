// decoder.cpp — scratch planner, staging and stage orchestration behind the C ABI.
// Replaces src/decoder.{hpp,cpp} + src/jpeggpu.cpp of the reference.  Differences by design:
//  * the scratch plan is computed once in parse_header (sizes are pure functions of the parsed
//    geometry), not by walking the pipeline twice with do_it=false/true;
//  * one packed staging block + one copy per scan (entropy-coded bytes only) instead of the
//    whole file plus 9+ table copies (src/decoder.cpp:175-208, quirk Q9);
//  * no memset (kernels clear what they need) and 7 launches per scan instead of ~40 stream
//    operations;
//  * cumulative scratch-size check (reference quirk Q7).
#include "plan.hpp"

#include <jpeggpu/jpeggpu.h>
#include <jpeggpu_b200.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

using namespace jb2;

struct jpeggpu_decoder {
    Logger log;
    ImagePlan plan;
    bool parsed = false;
    ScanLayout layout[kMaxScans];
    size_t upload_bytes = 0; // staging block size
    size_t total_bytes = 0;  // scratch requirement
    size_t transfer_bytes = 0;
    // Ring of pinned staging blocks.  parse_header fills a block, transfer enqueues its copy and
    // records an event behind it; the block is rewritten only after that event has completed,
    // so a caller may run several images ahead of the device with ONE decoder object (the
    // reference rewrites its pinned tables in place: SURVEY 8(b) "one image in flight").
    static constexpr int kStagingSlots = 4;
    struct StagingSlot {
        uint8_t* ptr = nullptr;
        size_t cap = 0;
        bool pinned = false;
        bool in_flight = false;
        bool have_event = false;
        int event_device = -1;
        cudaEvent_t copied{};
    } slots[kStagingSlots];
    int cur_slot = 0;
    uint8_t* staging = nullptr; // == slots[cur_slot].ptr
    LutSet lut_cache[kMaxScans];
    std::vector<uint8_t> lut_key;
    StageTimer timer;
    WalkPool* walk_pool = nullptr; // helper threads of the deferred walk (large scans only)
    bool walk_threads = true;

    ~jpeggpu_decoder()
    {
        walk_pool_destroy(walk_pool);
        for (StagingSlot& sl : slots) {
            // a copy out of the block may still be queued (jpeggpu_b200_decode_batch destroys its
            // lane decoders right after enqueueing): the pinned memory must outlive it
            if (sl.in_flight && sl.have_event) (void)cudaEventSynchronize(sl.copied);
            if (sl.have_event) cudaEventDestroy(sl.copied);
            release_slot(sl);
        }
        (void)cudaGetLastError();
        if (timer.created)
            for (int i = 0; i < StageTimer::kMaxMarks; ++i) cudaEventDestroy(timer.ev[i]);
    }

    static void release_slot(StagingSlot& sl)
    {
        if (sl.ptr) {
            if (sl.pinned) cudaFreeHost(sl.ptr);
            else free(sl.ptr);
        }
        sl.ptr = nullptr;
        sl.cap = 0;
    }

    // Pick the staging block for a freshly parsed image and make it at least `bytes` large.
    jpeggpu_status reserve_staging(size_t bytes)
    {
        if (slots[cur_slot].in_flight) cur_slot = (cur_slot + 1) % kStagingSlots;
        StagingSlot& sl = slots[cur_slot];
        if (sl.in_flight) { // the caller is kStagingSlots images ahead of the device: wait
            if (cudaEventSynchronize(sl.copied) != cudaSuccess) {
                (void)cudaGetLastError();
                return JPEGGPU_INTERNAL_ERROR;
            }
            sl.in_flight = false;
        }
        if (bytes > sl.cap) {
            release_slot(sl);
            const size_t cap = align256(bytes + bytes / 2 + 65536);
            void* p = nullptr;
            if (cudaHostAlloc(&p, cap, cudaHostAllocDefault) == cudaSuccess) {
                sl.pinned = true;
            } else {
                (void)cudaGetLastError(); // no device (CPU-only parse): pageable staging still works
                p = malloc(cap);
                sl.pinned = false;
                if (!p) return JPEGGPU_OUT_OF_HOST_MEMORY;
            }
            sl.ptr = static_cast<uint8_t*>(p);
            sl.cap = cap;
        }
        staging = sl.ptr;
        return JPEGGPU_SUCCESS;
    }

    // Called by transfer after the copy out of the current staging block has been enqueued.
    jpeggpu_status mark_staging_in_flight(cudaStream_t stream)
    {
        StagingSlot& sl = slots[cur_slot];
        if (!sl.pinned) return JPEGGPU_SUCCESS; // pageable source: the copy call already staged it
        if (record_staging_event(sl, stream)) {
            sl.in_flight = true;
            return JPEGGPU_SUCCESS;
        }
        // The copy is queued but cannot be guarded by an event: wait for it here, so that the
        // block is never rewritten under a pending DMA.
        (void)cudaGetLastError();
        return cudaStreamSynchronize(stream) == cudaSuccess ? JPEGGPU_SUCCESS : JPEGGPU_INTERNAL_ERROR;
    }

    bool record_staging_event(StagingSlot& sl, cudaStream_t stream)
    {
        int dev = -1;
        if (cudaGetDevice(&dev) != cudaSuccess) return false;
        if (sl.have_event && sl.event_device != dev) {
            cudaEventDestroy(sl.copied);
            sl.have_event = false;
        }
        if (!sl.have_event) {
            if (cudaEventCreateWithFlags(&sl.copied, cudaEventDisableTiming | cudaEventBlockingSync) != cudaSuccess)
                return false;
            sl.have_event = true;
            sl.event_device = dev;
        }
        return cudaEventRecord(sl.copied, stream) == cudaSuccess;
    }

    // Deferred entropy walk (parser.hpp): parse_header stops at the scan header of a single-scan
    // file and plans from upper bounds; transfer enqueues the upload of the scan bytes FIRST and
    // walks them on the host while the DMA runs, then uploads the directories.  Anything that
    // needs exact counts earlier (scan statistics) just runs the walk on the spot.
    int walk_mode = kWalkDevice; // WalkMode (parser.hpp)
    jpeggpu_status plan_scratch();
    void set_destuffed_bytes(int s);
    jpeggpu_status finish_walk();
    ScanBufs bufs(int s, void* d_tmp) const;
    jpeggpu_status rgbi_layout(int* cy, int* cb, int* cr) const;
    jpeggpu_status run(jpeggpu_img* img, int16_t* const* coef_planes, uint8_t* rgb, int rgb_pitch,
                       void* d_tmp, size_t tmp_size,
                       cudaStream_t stream);
};

// Lay out staging + scratch and fill the staging block (LUTs, slow tables, segment directory).
jpeggpu_status jpeggpu_decoder::plan_scratch()
{
    size_t off = 0;
    // pass 1: decode LUTs.  Expanding a DHT into the LUT pool costs tens of microseconds, and
    // consecutive images nearly always carry the same tables (same encoder): the pool of every
    // scan slot is cached and keyed by the raw table specifications.
    for (int s = 0; s < plan.num_scans; ++s) {
        ScanPlan& sp = plan.scans[s];
        LutSet& lc = lut_cache[s];
        lut_key_of(sp, lut_key);
        if (lc.key != lut_key) build_lut_set(sp, lut_key, lc);
        if (!apply_lut_set(lc, sp.geom)) return JPEGGPU_NOT_SUPPORTED; // > 160 KiB of LUTs
    }
    for (int s = 0; s < plan.num_scans; ++s) {
        ScanLayout& L = layout[s];
        const ScanGeom& g = plan.scans[s].geom;
        L.off_lut = off;
        off += align256((size_t)g.lut_pool_words * 4);
        L.off_slow = off;
        off += align256(sizeof(HuffSlow) * kMaxTables);
        L.off_desc = off;
        L.desc_words = destuff_desc_words(g);
        off += align256((size_t)L.desc_words * 4);
    }
    upload_bytes = off;
    transfer_bytes = upload_bytes;
    for (int s = 0; s < plan.num_scans; ++s) {
        layout[s].off_bytes = off;
        off += align256((size_t)plan.scans[s].geom.scan_avail + 16);
        transfer_bytes += plan.scans[s].geom.scan_avail;
    }
    // work area: sized for the largest scan
    const size_t work_begin = off;
    size_t work_end = off;
    for (int s = 0; s < plan.num_scans; ++s) {
        const size_t o = plan_work_area(plan.scans[s].geom, work_begin, layout[s]);
        set_destuffed_bytes(s);
        if (o > work_end) work_end = o;
    }
    total_bytes = work_end;

    jpeggpu_status st = reserve_staging(upload_bytes);
    if (st != JPEGGPU_SUCCESS) return st;
    for (int s = 0; s < plan.num_scans; ++s) {
        const ScanLayout& L = layout[s];
        const ScanPlan& sp = plan.scans[s];
        const LutSet& lc = lut_cache[s];
        memcpy(staging + L.off_lut, lc.pool.data(), (size_t)sp.geom.lut_pool_words * 4);
        memcpy(staging + L.off_slow, lc.slow, sizeof(HuffSlow) * (size_t)sp.geom.num_tables);
        memset(staging + L.off_desc, 0, (size_t)L.desc_words * 4);
    }
    return JPEGGPU_SUCCESS;
}

void jpeggpu_decoder::set_destuffed_bytes(int s)
{
    const ScanPlan& sp = plan.scans[s];
    layout[s].destuffed_bytes = sp.segs.empty() ? 0 : sp.segs.back().kept_before + sp.segs.back().nbytes;
}

// Runs the deferred walk (if one is pending): exact geometry into scans[0], counts re-derived
// inside the offsets planned from the bounds, directories into the staging block.
jpeggpu_status jpeggpu_decoder::finish_walk()
{
    if (!plan.walk_pending) return JPEGGPU_SUCCESS;
    const uint32_t planned_bytes = plan.scans[0].geom.scan_avail;
    const jpeggpu_status st = finish_deferred_walk(plan, log, walk_threads ? walk_pool : nullptr);
    if (st != JPEGGPU_SUCCESS) {
        parsed = false;
        return st;
    }
    transfer_bytes -= planned_bytes - plan.scans[0].geom.scan_avail;
    derive_counts(plan.scans[0].geom, layout[0]);
    set_destuffed_bytes(0);
    return JPEGGPU_SUCCESS;
}

ScanBufs jpeggpu_decoder::bufs(int s, void* d_tmp) const { return make_bufs(layout[s], d_tmp); }

#define JB2_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            log.log("CUDA error \"%s\" at %s:%d\n", cudaGetErrorString(e_), __FILE__,    \
                    __LINE__);                                                           \
            return JPEGGPU_INTERNAL_ERROR;                                               \
        }                                                                                \
    } while (0)

// The stage sequence of one image (reference decode_impl<true>, src/decoder.cpp:216-325).
// Which scan components carry Y / Cb / Cr when the image can go through idct_rgbi_kernel: one
// scan holding every frame component, grey or three components with 1x1 chroma (the layouts the
// reference's conv_to_rgbi accepts, util/util.h:43-69, minus chroma factors other than 1x1, which
// it mis-indexes).  *cb = -1 for grey.
jpeggpu_status jpeggpu_decoder::rgbi_layout(int* cy, int* cb, int* cr) const
{
    if (plan.num_comp != 1 && plan.num_comp != 3) return JPEGGPU_NOT_SUPPORTED;
    if (plan.num_scans != 1 || plan.scans[0].geom.num_comp != plan.num_comp) return JPEGGPU_NOT_SUPPORTED;
    const ScanGeom& g = plan.scans[0].geom;
    int at[3] = {-1, -1, -1};
    for (int a = 0; a < g.num_comp; ++a) at[g.frame_comp[a]] = a;
    *cy = at[0];
    *cb = *cr = -1;
    if (plan.num_comp == 1) return JPEGGPU_SUCCESS;
    if (at[0] < 0 || at[1] < 0 || at[2] < 0) return JPEGGPU_NOT_SUPPORTED;
    if (g.ssx[at[1]] != 1 || g.ssy[at[1]] != 1 || g.ssx[at[2]] != 1 || g.ssy[at[2]] != 1)
        return JPEGGPU_NOT_SUPPORTED;
    *cb = at[1];
    *cr = at[2];
    return JPEGGPU_SUCCESS;
}

jpeggpu_status jpeggpu_decoder::run(jpeggpu_img* img, int16_t* const* coef_planes, uint8_t* rgb,
                                    int rgb_pitch, void* d_tmp, size_t tmp_size, cudaStream_t stream)
{
    if (!parsed) return JPEGGPU_INVALID_ARGUMENT;
    if (plan.walk_pending && !plan.device_walk) return JPEGGPU_INVALID_ARGUMENT; // decode without transfer
    if (!d_tmp || tmp_size < total_bytes) return JPEGGPU_INTERNAL_ERROR; // src/util.hpp:66-68
    if ((reinterpret_cast<uintptr_t>(d_tmp) & 255u) != 0) return JPEGGPU_INVALID_ARGUMENT;
    timer.n = 0;
    for (int s = 0; s < plan.num_scans; ++s) {
        const ScanPlan& sp = plan.scans[s];
        const ScanGeom& g = sp.geom;
        const ScanBufs b = bufs(s, d_tmp);
        if (g.num_subs == 0) { // degenerate scan without entropy data: huff_sync will not run
            JB2_CUDA(cudaMemsetAsync(b.coef, 0, b.coef_bytes, stream));
            JB2_CUDA(cudaMemsetAsync(b.dc, 0, b.dc_bytes, stream));
        }
        if (g.scan_avail == 0) // destuff (which clears the control range) will not run either
            JB2_CUDA(cudaMemsetAsync(b.ctrl, 0, (size_t)b.clear_words * 4, stream));
        JB2_CUDA(launch_destuff(g, b, stream, &timer));
        JB2_CUDA(launch_huffman(g, b, stream, &timer));
        JB2_CUDA(launch_dc(g, b, stream, &timer));
        if (img) {
            OutPlanes out;
            memset(&out, 0, sizeof(out));
            for (int a = 0; a < g.num_comp; ++a) {
                const int ci = g.frame_comp[a];
                out.ptr[a] = img->image[ci];
                out.pitch[a] = img->pitch[ci];
                memcpy(out.q[a], plan.qtables[sp.qidx[a]], 64);
            }
            JB2_CUDA(launch_idct(g, b, out, stream, &timer));
        }
        if (rgb) {
            RgbOut out;
            memset(&out, 0, sizeof(out));
            const jpeggpu_status st = rgbi_layout(&out.cy, &out.cb, &out.cr);
            if (st != JPEGGPU_SUCCESS) return st;
            out.ptr = rgb;
            out.pitch = rgb_pitch;
            out.width = plan.width;
            out.height = plan.height;
            for (int a = 0; a < g.num_comp; ++a) memcpy(out.q[a], plan.qtables[sp.qidx[a]], 64);
            JB2_CUDA(launch_idct_rgbi(g, b, out, stream, &timer));
        }
        if (coef_planes) {
            CoefPlanes cp;
            memset(&cp, 0, sizeof(cp));
            for (int a = 0; a < g.num_comp; ++a) {
                cp.ptr[a] = coef_planes[g.frame_comp[a]];
                cp.stride[a] = g.blocks_x[a] * 8;
            }
            JB2_CUDA(launch_transpose(g, b, cp, stream, &timer));
        }
    }
    timer.mark(stream, kIdCount);
    return JPEGGPU_SUCCESS;
}

extern "C" {

// src/jpeggpu.cpp:41-60
const char* jpeggpu_get_status_string(enum jpeggpu_status stat)
{
    switch (stat) {
    case JPEGGPU_SUCCESS: return "success";
    case JPEGGPU_INVALID_ARGUMENT: return "invalid argument";
    case JPEGGPU_INVALID_JPEG: return "invalid jpeg";
    case JPEGGPU_INTERNAL_ERROR: return "internal jpeggpu error";
    case JPEGGPU_NOT_SUPPORTED: return "jpeg is not supported";
    case JPEGGPU_OUT_OF_HOST_MEMORY: return "out of host memory";
    case JPEGGPU_INCOMPLETE_BITSTREAM: return "incomplete bitstream";
    }
    return "unknown status";
}

// src/jpeggpu.cpp:62-75
enum jpeggpu_status jpeggpu_decoder_startup(jpeggpu_decoder_t* decoder)
{
    if (!decoder) return JPEGGPU_INVALID_ARGUMENT;
    *decoder = new (std::nothrow) jpeggpu_decoder();
    if (!*decoder) return JPEGGPU_OUT_OF_HOST_MEMORY;
    return JPEGGPU_SUCCESS;
}

// src/jpeggpu.cpp:77-86
enum jpeggpu_status jpeggpu_set_logging(jpeggpu_decoder_t decoder, int do_logging)
{
    if (!decoder) return JPEGGPU_INVALID_ARGUMENT;
    decoder->log.on = do_logging != 0;
    return JPEGGPU_SUCCESS;
}

// src/jpeggpu.cpp:88-104
int is_css_444(struct jpeggpu_subsampling css, int num_components)
{
    if (num_components < 1 || num_components > JPEGGPU_MAX_COMP) return 0;
    for (int c = 0; c < num_components; ++c)
        if (css.x[c] != 1 || css.y[c] != 1) return 0;
    return 1;
}

// src/jpeggpu.cpp:106-114, src/decoder.cpp:83-112
enum jpeggpu_status jpeggpu_decoder_parse_header(
    jpeggpu_decoder_t decoder, struct jpeggpu_img_info* img_info, const uint8_t* data, size_t size)
{
    if (!decoder || !img_info || !data) return JPEGGPU_INVALID_ARGUMENT;
    decoder->parsed = false;
#ifdef JB2_PARSE_TIMING
    const auto t0 = std::chrono::steady_clock::now();
#endif
    jpeggpu_status st = parse_jpeg(data, size, decoder->plan, decoder->log, decoder->walk_mode);
    if (st != JPEGGPU_SUCCESS) return st;
#ifdef JB2_PARSE_TIMING
    const auto t1 = std::chrono::steady_clock::now();
#endif
    try {
        st = decoder->plan_scratch();
#ifdef JB2_PARSE_TIMING
        const auto t2 = std::chrono::steady_clock::now();
        fprintf(stderr, "parse_jpeg %.1f us plan_scratch %.1f us\n",
                std::chrono::duration<double, std::micro>(t1 - t0).count(),
                std::chrono::duration<double, std::micro>(t2 - t1).count());
#endif
    } catch (const std::bad_alloc&) {
        st = JPEGGPU_OUT_OF_HOST_MEMORY;
    }
    if (st != JPEGGPU_SUCCESS) return st;
    const ImagePlan& p = decoder->plan;
    memset(img_info, 0, sizeof(*img_info));
    img_info->num_components = p.num_comp;
    for (int c = 0; c < p.num_comp; ++c) {
        img_info->sizes_x[c] = p.comps[c].size_x;
        img_info->sizes_y[c] = p.comps[c].size_y;
        img_info->subsampling.x[c] = p.comps[c].ssx;
        img_info->subsampling.y[c] = p.comps[c].ssy;
    }
    decoder->parsed = true;
    return JPEGGPU_SUCCESS;
}

// src/jpeggpu.cpp:116-123
enum jpeggpu_status jpeggpu_decoder_get_buffer_size(jpeggpu_decoder_t decoder, size_t* tmp_size)
{
    if (!decoder || !tmp_size) return JPEGGPU_INVALID_ARGUMENT;
    if (!decoder->parsed) return JPEGGPU_INVALID_ARGUMENT;
    *tmp_size = decoder->total_bytes;
    return JPEGGPU_SUCCESS;
}

// src/jpeggpu.cpp:125-133, src/decoder.cpp:159-211
enum jpeggpu_status jpeggpu_decoder_transfer(
    jpeggpu_decoder_t decoder, void* d_tmp, size_t tmp_size, cudaStream_t stream)
{
    if (!decoder || !decoder->parsed) return JPEGGPU_INVALID_ARGUMENT;
    if (!d_tmp || tmp_size < decoder->total_bytes) return JPEGGPU_INTERNAL_ERROR;
    const Logger& log = decoder->log;
    uint8_t* base = static_cast<uint8_t*>(d_tmp);
    if (decoder->plan.walk_pending && !decoder->plan.device_walk) {
        // the big copy goes first; the host walks the same bytes while the DMA engine reads them
        const ScanPlan& sp = decoder->plan.scans[0];
        if (decoder->walk_threads && decoder->plan.file_size - sp.begin >= kWalkParallelMin) {
            // wake the helpers before the enqueue below so that their start-up hides behind it
            if (!decoder->walk_pool) decoder->walk_pool = walk_pool_create();
            walk_pool_kick(decoder->walk_pool, decoder->plan);
        }
        if (cudaMemcpyAsync(base + decoder->layout[0].off_bytes, decoder->plan.file + sp.begin,
                            sp.geom.scan_avail, cudaMemcpyHostToDevice, stream) != cudaSuccess) {
            (void)cudaGetLastError();
            walk_pool_cancel(decoder->walk_pool); // no helper may read the caller's buffer after we return
            decoder->parsed = false;
            return JPEGGPU_INTERNAL_ERROR;
        }
        const jpeggpu_status st = decoder->finish_walk();
        if (st != JPEGGPU_SUCCESS) return st;
        JB2_CUDA(cudaMemcpyAsync(base, decoder->staging, decoder->upload_bytes,
                                 cudaMemcpyHostToDevice, stream));
        if (decoder->mark_staging_in_flight(stream) != JPEGGPU_SUCCESS) return JPEGGPU_INTERNAL_ERROR;
        return JPEGGPU_SUCCESS;
    }
    JB2_CUDA(cudaMemcpyAsync(base, decoder->staging, decoder->upload_bytes,
                             cudaMemcpyHostToDevice, stream));
    if (decoder->mark_staging_in_flight(stream) != JPEGGPU_SUCCESS) return JPEGGPU_INTERNAL_ERROR;
    for (int s = 0; s < decoder->plan.num_scans; ++s) {
        const ScanPlan& sp = decoder->plan.scans[s];
        if (sp.geom.scan_avail == 0) continue;
        JB2_CUDA(cudaMemcpyAsync(base + decoder->layout[s].off_bytes, decoder->plan.file + sp.begin,
                                 sp.geom.scan_avail, cudaMemcpyHostToDevice, stream));
    }
    return JPEGGPU_SUCCESS;
}

// src/jpeggpu.cpp:135-147, src/decoder.cpp:336-354
enum jpeggpu_status jpeggpu_decoder_decode(
    jpeggpu_decoder_t decoder, struct jpeggpu_img* img, void* d_tmp, size_t tmp_size,
    cudaStream_t stream)
{
    if (!decoder || !img) return JPEGGPU_INVALID_ARGUMENT;
    if (!decoder->parsed) return JPEGGPU_INVALID_ARGUMENT;
    const ImagePlan& p = decoder->plan;
    for (int c = 0; c < p.num_comp; ++c)
        if (!img->image[c] || img->pitch[c] < p.comps[c].size_x) return JPEGGPU_INVALID_ARGUMENT;
    return decoder->run(img, nullptr, nullptr, 0, d_tmp, tmp_size, stream);
}

// src/jpeggpu.cpp:149-159
enum jpeggpu_status jpeggpu_decoder_cleanup(jpeggpu_decoder_t decoder)
{
    if (!decoder) return JPEGGPU_INVALID_ARGUMENT;
    delete decoder;
    return JPEGGPU_SUCCESS;
}

// ---- additive entry points (include/jpeggpu_b200.h) --------------------------------------------

static const ScanGeom* coding_scan(const ImagePlan& p, int comp, int* a_out)
{
    for (int s = 0; s < p.num_scans; ++s)
        for (int a = 0; a < p.scans[s].geom.num_comp; ++a)
            if (p.scans[s].geom.frame_comp[a] == comp) {
                *a_out = a;
                return &p.scans[s].geom;
            }
    return nullptr;
}

enum jpeggpu_status jpeggpu_b200_get_coefficient_size(
    jpeggpu_decoder_t decoder, int width[JPEGGPU_MAX_COMP], int height[JPEGGPU_MAX_COMP])
{
    if (!decoder || !width || !height || !decoder->parsed) return JPEGGPU_INVALID_ARGUMENT;
    for (int c = 0; c < JPEGGPU_MAX_COMP; ++c) {
        width[c] = height[c] = 0;
        int a = 0;
        const ScanGeom* g = c < decoder->plan.num_comp ? coding_scan(decoder->plan, c, &a) : nullptr;
        if (g) {
            width[c] = g->blocks_x[a] * 8;
            height[c] = g->blocks_y[a] * 8;
        }
    }
    return JPEGGPU_SUCCESS;
}

enum jpeggpu_status jpeggpu_b200_decode_coefficients(
    jpeggpu_decoder_t decoder, int16_t* d_planes[JPEGGPU_MAX_COMP], void* d_tmp, size_t tmp_size,
    cudaStream_t stream)
{
    if (!decoder || !d_planes || !decoder->parsed) return JPEGGPU_INVALID_ARGUMENT;
    for (int c = 0; c < decoder->plan.num_comp; ++c)
        if (!d_planes[c] || (reinterpret_cast<uintptr_t>(d_planes[c]) & 15u))
            return JPEGGPU_INVALID_ARGUMENT;
    return decoder->run(nullptr, d_planes, nullptr, 0, d_tmp, tmp_size, stream);
}

enum jpeggpu_status jpeggpu_b200_decode_rgbi(
    jpeggpu_decoder_t decoder, uint8_t* d_rgb, int pitch, void* d_tmp, size_t tmp_size,
    cudaStream_t stream)
{
    if (!decoder || !d_rgb || !decoder->parsed) return JPEGGPU_INVALID_ARGUMENT;
    if ((long long)pitch < 3ll * decoder->plan.width) return JPEGGPU_INVALID_ARGUMENT;
    int cy, cb, cr;
    const jpeggpu_status st = decoder->rgbi_layout(&cy, &cb, &cr);
    if (st != JPEGGPU_SUCCESS) return st;
    return decoder->run(nullptr, nullptr, d_rgb, pitch, d_tmp, tmp_size, stream);
}

enum jpeggpu_status jpeggpu_b200_convert_rgbi(
    const struct jpeggpu_img* planes, const struct jpeggpu_img_info* info, uint8_t* d_rgb,
    int pitch, cudaStream_t stream)
{
    if (!planes || !info || !d_rgb) return JPEGGPU_INVALID_ARGUMENT;
    const int nc = info->num_components;
    if (nc != 1 && nc != 3) return JPEGGPU_NOT_SUPPORTED; // util/util.h:43-46
    const jpeggpu_subsampling& ss = info->subsampling;
    PlanarIn in;
    memset(&in, 0, sizeof(in));
    in.num_comp = nc;
    in.ssx = in.ssy = 1;
    if (nc == 3) {
        // util/util.h:61-69 refuses subsampled luma and unequal chroma; chroma factors other
        // than 1x1 (which it would mis-index) are refused here as well
        if (ss.x[1] != 1 || ss.y[1] != 1 || ss.x[2] != 1 || ss.y[2] != 1) return JPEGGPU_NOT_SUPPORTED;
        if (ss.x[0] < 1 || ss.x[0] > 4 || ss.y[0] < 1 || ss.y[0] > 4) return JPEGGPU_INVALID_ARGUMENT;
        in.ssx = ss.x[0];
        in.ssy = ss.y[0];
    }
    const int w = info->sizes_x[0], h = info->sizes_y[0];
    if (w < 0 || h < 0 || (long long)pitch < 3ll * w) return JPEGGPU_INVALID_ARGUMENT;
    for (int c = 0; c < nc; ++c) {
        if (!planes->image[c] || planes->pitch[c] < info->sizes_x[c]) return JPEGGPU_INVALID_ARGUMENT;
        in.ptr[c] = planes->image[c];
        in.pitch[c] = planes->pitch[c];
    }
    RgbOut out;
    memset(&out, 0, sizeof(out));
    out.ptr = d_rgb;
    out.pitch = pitch;
    out.width = w;
    out.height = h;
    if (launch_ycc_to_rgbi(in, out, stream) != cudaSuccess) {
        (void)cudaGetLastError();
        return JPEGGPU_INTERNAL_ERROR;
    }
    return JPEGGPU_SUCCESS;
}

enum jpeggpu_status jpeggpu_b200_get_scan_stats(
    jpeggpu_decoder_t decoder, int scan, struct jpeggpu_b200_scan_stats* stats, int* num_scans)
{
    if (!decoder || !decoder->parsed) return JPEGGPU_INVALID_ARGUMENT;
    {
        const jpeggpu_status st = decoder->finish_walk(); // exact counts need the walk
        if (st != JPEGGPU_SUCCESS) return st;
    }
    if (num_scans) *num_scans = decoder->plan.num_scans;
    if (!stats) return JPEGGPU_SUCCESS;
    if (scan < 0 || scan >= decoder->plan.num_scans) return JPEGGPU_INVALID_ARGUMENT;
    const ScanGeom& g = decoder->plan.scans[scan].geom;
    const ScanLayout& L = decoder->layout[scan];
    memset(stats, 0, sizeof(*stats));
    stats->num_components = g.num_comp;
    stats->num_segments = g.num_segments;
    stats->dus_per_mcu = g.dus_per_mcu;
    stats->total_mcus = g.total_mcus;
    stats->scan_bytes = g.scan_bytes;
    stats->destuffed_bytes = L.destuffed_bytes;
    stats->num_subsequences = g.num_subs;
    stats->num_data_units = L.num_dus;
    stats->subsequence_bytes = kSubBytes;
    stats->kernel_launches = count_launches(g, false);
    stats->off_control = L.off_ctrl;
    stats->off_destuffed = L.off_destuffed;
    stats->off_coefficients = L.off_coef;
    stats->off_sync_records = L.off_entries;
    stats->off_slot_prefix = L.off_qpos;
    stats->off_segments = L.off_segs;
    stats->off_scan_info = L.off_info;
    return JPEGGPU_SUCCESS;
}

enum jpeggpu_status jpeggpu_b200_read_scan_info(
    jpeggpu_decoder_t decoder, int scan, const void* d_tmp, cudaStream_t stream,
    struct jpeggpu_b200_scan_info* info)
{
    if (!decoder || !decoder->parsed || !d_tmp || !info) return JPEGGPU_INVALID_ARGUMENT;
    if (scan < 0 || scan >= decoder->plan.num_scans) return JPEGGPU_INVALID_ARGUMENT;
    static_assert(sizeof(ScanInfo) == sizeof(jpeggpu_b200_scan_info), "same four words");
    const uint8_t* src = static_cast<const uint8_t*>(d_tmp) + decoder->layout[scan].off_info;
    if (cudaMemcpyAsync(info, src, sizeof(*info), cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
        cudaStreamSynchronize(stream) != cudaSuccess) {
        (void)cudaGetLastError();
        return JPEGGPU_INTERNAL_ERROR;
    }
    return JPEGGPU_SUCCESS;
}

enum jpeggpu_status jpeggpu_b200_get_transfer_size(jpeggpu_decoder_t decoder, size_t* bytes)
{
    if (!decoder || !bytes || !decoder->parsed) return JPEGGPU_INVALID_ARGUMENT;
    {
        const jpeggpu_status st = decoder->finish_walk();
        if (st != JPEGGPU_SUCCESS) return st;
    }
    *bytes = decoder->transfer_bytes;
    return JPEGGPU_SUCCESS;
}

enum jpeggpu_status jpeggpu_b200_set_deferred_walk(jpeggpu_decoder_t decoder, int on)
{
    if (!decoder) return JPEGGPU_INVALID_ARGUMENT;
    if (on < 0 || on > 3) return JPEGGPU_INVALID_ARGUMENT;
    decoder->walk_mode = on == 0 ? kWalkEager : on == 3 ? kWalkDevice : kWalkDeferred;
    decoder->walk_threads = on != 2;
    return JPEGGPU_SUCCESS;
}

enum jpeggpu_status jpeggpu_b200_set_profiling(jpeggpu_decoder_t decoder, int on)
{
    if (!decoder) return JPEGGPU_INVALID_ARGUMENT;
    StageTimer& t = decoder->timer;
    if (on && !t.created) {
        for (int i = 0; i < StageTimer::kMaxMarks; ++i)
            if (cudaEventCreate(&t.ev[i]) != cudaSuccess) return JPEGGPU_INTERNAL_ERROR;
        t.created = true;
    }
    t.on = on != 0;
    t.n = 0;
    return JPEGGPU_SUCCESS;
}

enum jpeggpu_status jpeggpu_b200_get_kernel_times(
    jpeggpu_decoder_t decoder, float ms[JPEGGPU_B200_MAX_KERNELS], int launches[JPEGGPU_B200_MAX_KERNELS])
{
    if (!decoder || !ms) return JPEGGPU_INVALID_ARGUMENT;
    StageTimer& t = decoder->timer;
    for (int i = 0; i < JPEGGPU_B200_MAX_KERNELS; ++i) {
        ms[i] = 0.f;
        if (launches) launches[i] = 0;
    }
    if (!t.on) return JPEGGPU_INVALID_ARGUMENT;
    for (int i = 0; i + 1 < t.n; ++i) {
        float d = 0.f;
        if (cudaEventElapsedTime(&d, t.ev[i], t.ev[i + 1]) != cudaSuccess)
            return JPEGGPU_INTERNAL_ERROR; // the caller has not synchronised the stream yet
        if (t.id[i] < JPEGGPU_B200_MAX_KERNELS) {
            ms[t.id[i]] += d;
            if (launches) launches[t.id[i]] += 1;
        }
    }
    return JPEGGPU_SUCCESS;
}

const char* jpeggpu_b200_kernel_name(int id)
{
    static const char* names[kIdCount] = {
        "destuff_kernel", "huff_sync", "huff_write", "dc_scan",
        "idct_kernel", "transpose_kernel", "idct_rgbi_kernel"};
    return id >= 0 && id < kIdCount ? names[id] : "";
}

} // extern "C"
