// batch.cpp — native driver for batches of independent images (include/jpeggpu_b200.h).
// It is a caller of the per-image C API, exactly like the loop in the reference's benchmark
// (benchmark/benchmark_jpeggpu.hpp:69-89), run by one host thread per stream.
#include <jpeggpu/jpeggpu.h>
#include <jpeggpu_b200.h>

#include <cuda_runtime.h>

#include <atomic>
#include <new>
#include <thread>
#include <vector>

namespace {

struct DecoderGuard {
    jpeggpu_decoder_t d = nullptr;
    jpeggpu_status start() { return jpeggpu_decoder_startup(&d); }
    ~DecoderGuard()
    {
        if (d) jpeggpu_decoder_cleanup(d);
    }
};

jpeggpu_status decode_one(jpeggpu_decoder_t dec, jpeggpu_b200_batch_item& it, void* d_tmp,
                          size_t tmp_size, cudaStream_t stream)
{
    jpeggpu_status st = jpeggpu_decoder_parse_header(dec, &it.info, it.data, it.size);
    if (st != JPEGGPU_SUCCESS) return st;
    size_t need = 0;
    st = jpeggpu_decoder_get_buffer_size(dec, &need);
    if (st != JPEGGPU_SUCCESS) return st;
    if (need > tmp_size) return JPEGGPU_INTERNAL_ERROR;
    st = jpeggpu_decoder_transfer(dec, d_tmp, tmp_size, stream);
    if (st != JPEGGPU_SUCCESS) return st;
    return jpeggpu_decoder_decode(dec, &it.img, d_tmp, tmp_size, stream);
}

void lane_main(int device, jpeggpu_b200_batch_item* items, int num_items, int lane, int num_lanes,
               cudaStream_t stream, void* d_tmp, size_t tmp_size)
{
    jpeggpu_status fail = JPEGGPU_SUCCESS;
    DecoderGuard g;
    if (cudaSetDevice(device) != cudaSuccess) fail = JPEGGPU_INTERNAL_ERROR;
    if (fail == JPEGGPU_SUCCESS) fail = g.start();
    for (int i = lane; i < num_items; i += num_lanes) {
        jpeggpu_b200_batch_item& it = items[i];
        if (fail != JPEGGPU_SUCCESS) it.status = fail;
        else if (!it.data) it.status = JPEGGPU_INVALID_ARGUMENT;
        else it.status = decode_one(g.d, it, d_tmp, tmp_size, stream);
    }
}

template <class Fn>
jpeggpu_status run_threads(int n, Fn&& body)
{
    try {
        std::vector<std::thread> th;
        th.reserve((size_t)n > 0 ? (size_t)n - 1 : 0);
        for (int t = 1; t < n; ++t) th.emplace_back(body, t);
        body(0); // the calling thread is lane 0
        for (std::thread& t : th) t.join();
    } catch (const std::bad_alloc&) {
        return JPEGGPU_OUT_OF_HOST_MEMORY;
    } catch (...) {
        return JPEGGPU_INTERNAL_ERROR;
    }
    return JPEGGPU_SUCCESS;
}

jpeggpu_status first_failure(const jpeggpu_b200_batch_item* items, int n)
{
    for (int i = 0; i < n; ++i)
        if (items[i].status != JPEGGPU_SUCCESS) return items[i].status;
    return JPEGGPU_SUCCESS;
}

} // namespace

extern "C" {

enum jpeggpu_status jpeggpu_b200_decode_batch(
    struct jpeggpu_b200_batch_item* items, int num_items, const cudaStream_t* streams,
    void* const* d_tmp, const size_t* tmp_size, int num_lanes)
{
    if (!items || num_items < 0 || !streams || !d_tmp || !tmp_size || num_lanes < 1)
        return JPEGGPU_INVALID_ARGUMENT;
    for (int l = 0; l < num_lanes; ++l)
        if (!d_tmp[l]) return JPEGGPU_INVALID_ARGUMENT;
    int device = 0;
    if (cudaGetDevice(&device) != cudaSuccess) return JPEGGPU_INTERNAL_ERROR;
    if (num_lanes > num_items) num_lanes = num_items > 0 ? num_items : 1;
    const jpeggpu_status st = run_threads(num_lanes, [=](int lane) {
        lane_main(device, items, num_items, lane, num_lanes, streams[lane], d_tmp[lane], tmp_size[lane]);
    });
    if (st != JPEGGPU_SUCCESS) return st;
    return first_failure(items, num_items);
}

enum jpeggpu_status jpeggpu_b200_parse_batch(
    struct jpeggpu_b200_batch_item* items, int num_items, int num_threads, size_t* max_tmp_size)
{
    if (!items || num_items < 0 || num_threads < 1) return JPEGGPU_INVALID_ARGUMENT;
    if (num_threads > num_items) num_threads = num_items > 0 ? num_items : 1;
    std::atomic<size_t> biggest{0};
    const jpeggpu_status st = run_threads(num_threads, [&, num_threads](int t) {
        DecoderGuard g;
        const jpeggpu_status start = g.start();
        for (int i = t; i < num_items; i += num_threads) {
            jpeggpu_b200_batch_item& it = items[i];
            if (start != JPEGGPU_SUCCESS) {
                it.status = start;
                continue;
            }
            it.status = it.data ? jpeggpu_decoder_parse_header(g.d, &it.info, it.data, it.size)
                                : JPEGGPU_INVALID_ARGUMENT;
            size_t need = 0;
            if (it.status == JPEGGPU_SUCCESS &&
                jpeggpu_decoder_get_buffer_size(g.d, &need) == JPEGGPU_SUCCESS) {
                size_t seen = biggest.load();
                while (need > seen && !biggest.compare_exchange_weak(seen, need)) {
                }
            }
        }
    });
    if (max_tmp_size) *max_tmp_size = biggest.load();
    if (st != JPEGGPU_SUCCESS) return st;
    return first_failure(items, num_items);
}

} // extern "C"

// ---- fused batches (include/jpeggpu_b200.h): several images per kernel launch ------------------
#include "plan.hpp"

#ifndef JB2_BATCH_FILL_IN_WRITE
#define JB2_BATCH_FILL_IN_WRITE 1
#endif

#include <algorithm>
#include <cstring>
#include <memory>

struct jpeggpu_b200_batch {
    struct Image {
        jb2::ImagePlan plan;
        jb2::ScanLayout layout;
        int lut_set = -1;
        uint32_t lead = 0;
        bool ok = false;
    };
    std::vector<std::unique_ptr<Image>> images; // reused across batches (the plans keep their vectors)
    std::vector<std::unique_ptr<jb2::LutSet>> lut_sets; // table sets seen so far (a handful in practice)
    std::vector<size_t> lut_off, slow_off;      // arena offsets of the sets used by the current batch
    std::vector<int> used_sets;
    std::vector<uint8_t> key;
    jb2::Logger log;
    int planned = 0;
    size_t off_desc = 0, upload_bytes = 0, zero_begin = 0, zero_bytes = 0, arena_bytes = 0;
    // file bytes: one span copy (only inside a host range the caller vouched for), or one per file
    const uint8_t* host_lo = nullptr;
    const uint8_t* host_hi = nullptr;
    bool span_copy = false;
    const uint8_t* span_src = nullptr;
    size_t span_bytes = 0, off_span = 0;
    uint32_t max_pool_words = 0, max_compact_words = 0;
    // pinned staging block of the descriptor + table upload, guarded by an event
    uint8_t* staging = nullptr;
    size_t staging_cap = 0;
    bool pinned = false, in_flight = false, have_event = false;
    int event_device = -1;
    cudaEvent_t copied{};

    ~jpeggpu_b200_batch()
    {
        if (in_flight && have_event) (void)cudaEventSynchronize(copied);
        if (have_event) cudaEventDestroy(copied);
        if (staging) {
            if (pinned) cudaFreeHost(staging);
            else free(staging);
        }
        (void)cudaGetLastError();
    }

    jpeggpu_status reserve(size_t bytes)
    {
        if (in_flight) {
            if (cudaEventSynchronize(copied) != cudaSuccess) {
                (void)cudaGetLastError();
                return JPEGGPU_INTERNAL_ERROR;
            }
            in_flight = false;
        }
        if (bytes <= staging_cap) return JPEGGPU_SUCCESS;
        if (staging) {
            if (pinned) cudaFreeHost(staging);
            else free(staging);
            staging = nullptr;
        }
        const size_t cap = jb2::align256(bytes + bytes / 2 + 65536);
        void* p = nullptr;
        if (cudaHostAlloc(&p, cap, cudaHostAllocDefault) == cudaSuccess) {
            pinned = true;
        } else {
            (void)cudaGetLastError(); // no device (host-only planning): pageable staging still works
            p = malloc(cap);
            pinned = false;
            if (!p) return JPEGGPU_OUT_OF_HOST_MEMORY;
        }
        staging = static_cast<uint8_t*>(p);
        staging_cap = cap;
        return JPEGGPU_SUCCESS;
    }
};

namespace {

using namespace jb2;

jpeggpu_status plan_batch(jpeggpu_b200_batch& B, jpeggpu_b200_batch_item* items, int n, size_t* arena_size)
{
    while ((int)B.images.size() < n) B.images.emplace_back(new jpeggpu_b200_batch::Image);
    if (B.lut_sets.size() > 64) B.lut_sets.clear(); // a stream of ever new tables: start over
    B.planned = 0;
    B.used_sets.clear();
    B.max_pool_words = B.max_compact_words = 0;
    jpeggpu_status first = JPEGGPU_SUCCESS;
    // 1. headers, table sets
    const uint8_t* lo = nullptr;
    const uint8_t* hi = nullptr;
    size_t sum_sizes = 0;
    int last_set = -1;
    for (int i = 0; i < n; ++i) {
        jpeggpu_b200_batch::Image& im = *B.images[i];
        jpeggpu_b200_batch_item& it = items[i];
        im.ok = false;
        memset(&it.info, 0, sizeof(it.info));
        jpeggpu_status st = it.data ? parse_jpeg(it.data, it.size, im.plan, B.log, kWalkDeviceAlways)
                                    : JPEGGPU_INVALID_ARGUMENT;
        if (st == JPEGGPU_SUCCESS && !(im.plan.walk_pending && im.plan.device_walk))
            st = JPEGGPU_NOT_SUPPORTED; // several scans: the per-image path decodes those
        if (st == JPEGGPU_SUCCESS) {
            ScanPlan& sp = im.plan.scans[0];
            lut_key_of(sp, B.key);
            int set = -1;
            if (last_set >= 0 && B.lut_sets[last_set]->key == B.key) set = last_set;
            for (int k = 0; set < 0 && k < (int)B.lut_sets.size(); ++k)
                if (B.lut_sets[k]->key == B.key) set = k;
            if (set < 0) {
                B.lut_sets.emplace_back(new LutSet);
                set = (int)B.lut_sets.size() - 1;
                build_lut_set(sp, B.key, *B.lut_sets[set]);
            }
            last_set = set;
            if (!apply_lut_set(*B.lut_sets[set], sp.geom)) st = JPEGGPU_NOT_SUPPORTED;
            im.lut_set = set;
        }
        it.status = st;
        if (st != JPEGGPU_SUCCESS) {
            if (first == JPEGGPU_SUCCESS) first = st;
            continue;
        }
        im.ok = true;
        if (std::find(B.used_sets.begin(), B.used_sets.end(), im.lut_set) == B.used_sets.end())
            B.used_sets.push_back(im.lut_set);
        const ImagePlan& p = im.plan;
        it.info.num_components = p.num_comp;
        for (int c = 0; c < p.num_comp; ++c) {
            it.info.sizes_x[c] = p.comps[c].size_x;
            it.info.sizes_y[c] = p.comps[c].size_y;
            it.info.subsampling.x[c] = p.comps[c].ssx;
            it.info.subsampling.y[c] = p.comps[c].ssy;
        }
        const ScanGeom& g = p.scans[0].geom;
        B.max_pool_words = std::max(B.max_pool_words, g.lut_pool_words);
        B.max_compact_words = std::max(B.max_compact_words, g.lut_compact_words);
        const uint8_t* a = it.data + p.scans[0].begin;
        const uint8_t* e = it.data + it.size;
        lo = !lo || a < lo ? a : lo;
        hi = !hi || e > hi ? e : hi;
        sum_sizes += (size_t)(e - a);
    }
    // 2. arena: [descriptors | table sets] [zeros: look-back descriptors of destuff] [file bytes] [work areas]
    size_t off = 0;
    B.off_desc = off;
    off += align256(sizeof(BatchImage) * (size_t)n);
    B.lut_off.assign(B.lut_sets.size(), 0);
    B.slow_off.assign(B.lut_sets.size(), 0);
    for (int k : B.used_sets) {
        B.lut_off[k] = off;
        off += align256(B.lut_sets[k]->pool.size() * 4);
        B.slow_off[k] = off;
        off += align256(sizeof(HuffSlow) * kMaxTables);
    }
    B.upload_bytes = off;
    B.zero_begin = off;
    for (int i = 0; i < n; ++i) {
        jpeggpu_b200_batch::Image& im = *B.images[i];
        if (!im.ok) continue;
        const ScanGeom& g = im.plan.scans[0].geom;
        im.layout = ScanLayout{};
        im.layout.off_lut = B.lut_off[im.lut_set];
        im.layout.off_slow = B.slow_off[im.lut_set];
        im.layout.off_desc = off;
        im.layout.desc_words = destuff_desc_words(g);
        off += ((size_t)im.layout.desc_words * 4 + 15) & ~(size_t)15;
    }
    off = align256(off);
    B.zero_bytes = off - B.zero_begin;
    // the files lie close together in host memory: one copy of the whole span (headers included)
    B.span_copy = false;
    if (lo) {
        const uint8_t* base = reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(lo) & ~(uintptr_t)15);
        const size_t span = (size_t)(hi - base);
        if (B.host_lo && base >= B.host_lo && hi <= B.host_hi && span <= sum_sizes + sum_sizes / 2 + 65536 &&
            span < (1ull << 31)) {
            B.span_copy = true;
            B.span_src = base;
            B.span_bytes = span;
            B.off_span = off;
            off += align256(span + 16);
        }
    }
    for (int i = 0; i < n; ++i) {
        jpeggpu_b200_batch::Image& im = *B.images[i];
        if (!im.ok) continue;
        const ScanGeom& g = im.plan.scans[0].geom;
        if (B.span_copy) {
            const size_t rel = (size_t)(items[i].data + im.plan.scans[0].begin - B.span_src);
            im.layout.off_bytes = B.off_span + (rel & ~(size_t)15);
            im.lead = (uint32_t)(rel & 15);
        } else {
            im.layout.off_bytes = off;
            im.lead = 0;
            off += align256((size_t)g.scan_avail + 16);
        }
    }
    for (int i = 0; i < n; ++i) {
        jpeggpu_b200_batch::Image& im = *B.images[i];
        if (im.ok) off = plan_work_area(im.plan.scans[0].geom, off, im.layout);
    }
    B.arena_bytes = off;
    B.planned = n;
    if (arena_size) *arena_size = off;
    // 3. the table sets into the staging block (the descriptors follow at decode time: they hold
    //    device addresses)
    const jpeggpu_status rs = B.reserve(B.upload_bytes);
    if (rs != JPEGGPU_SUCCESS) return rs;
    for (int k : B.used_sets) {
        const LutSet& ls = *B.lut_sets[k];
        memcpy(B.staging + B.lut_off[k], ls.pool.data(), ls.pool.size() * 4);
        memcpy(B.staging + B.slow_off[k], ls.slow, sizeof(ls.slow));
    }
    return first;
}

jpeggpu_status decode_batch(jpeggpu_b200_batch& B, jpeggpu_b200_batch_item* items, int n, void* d_arena,
                            size_t arena_size, cudaStream_t stream)
{
    if (n != B.planned) return JPEGGPU_INVALID_ARGUMENT;
    if (arena_size < B.arena_bytes) return JPEGGPU_INTERNAL_ERROR; // src/util.hpp:66-68
    if ((reinterpret_cast<uintptr_t>(d_arena) & 255u) != 0) return JPEGGPU_INVALID_ARGUMENT;
    uint8_t* base = static_cast<uint8_t*>(d_arena);
    BatchImage* desc = reinterpret_cast<BatchImage*>(B.staging + B.off_desc);
    BatchLaunch bl;
    memset(&bl, 0, sizeof(bl));
    bl.d_images = reinterpret_cast<const BatchImage*>(base + B.off_desc);
    bl.count = (uint32_t)n;
    bl.lut_pool_words_max = B.max_pool_words;
    bl.lut_compact_words_max = B.max_compact_words;
    const int write_warps = batch_write_warps(B.max_compact_words);
    jpeggpu_status first = JPEGGPU_SUCCESS;
    size_t batch_coef_bytes = 0;
    for (int i = 0; i < n; ++i)
        if (B.images[i]->ok) batch_coef_bytes += B.images[i]->layout.coef_bytes;
    for (int i = 0; i < n; ++i) {
        jpeggpu_b200_batch::Image& im = *B.images[i];
        BatchImage& d = desc[i];
        memset(&d, 0, sizeof(d)); // ctas == 0: every kernel skips the image
        if (!im.ok) continue;
        const ImagePlan& p = im.plan;
        const ScanPlan& sp = p.scans[0];
        bool good = true;
        for (int c = 0; c < p.num_comp; ++c)
            good &= items[i].img.image[c] != nullptr && items[i].img.pitch[c] >= p.comps[c].size_x;
        if (!good) {
            items[i].status = JPEGGPU_INVALID_ARGUMENT;
            if (first == JPEGGPU_SUCCESS) first = JPEGGPU_INVALID_ARGUMENT;
            continue;
        }
        d.g = sp.geom;
        d.b = make_bufs(im.layout, base);
        // the coefficient buffers of a batch are written and read together: what counts for "fits
        // L2" is their sum (64 files of 2 MP: 397 MB)
        if (JB2_BATCH_FILL_IN_WRITE && batch_coef_bytes > kFillInWriteBytes) d.b.fill_in_write = 1u;
        d.b.scan_lead = im.lead;
        uint8_t q[kMaxComp][64];
        memset(q, 0, sizeof(q));
        for (int a = 0; a < sp.geom.num_comp; ++a) {
            const int ci = sp.geom.frame_comp[a];
            d.plane[a] = items[i].img.image[ci];
            d.pitch[a] = items[i].img.pitch[ci];
            memcpy(q[a], p.qtables[sp.qidx[a]], 64);
        }
        batch_image_grid(d, q, write_warps);
        for (int k = 0; k < kBkCount; ++k) bl.max_ctas[k] = std::max(bl.max_ctas[k], d.ctas[k]);
    }
    if (bl.max_ctas[kBkDestuff] == 0) return first; // nothing to decode
    if (cudaMemcpyAsync(base, B.staging, B.upload_bytes, cudaMemcpyHostToDevice, stream) != cudaSuccess ||
        cudaMemsetAsync(base + B.zero_begin, 0, B.zero_bytes, stream) != cudaSuccess) {
        (void)cudaGetLastError();
        return JPEGGPU_INTERNAL_ERROR;
    }
    if (B.span_copy) {
        if (cudaMemcpyAsync(base + B.off_span, B.span_src, B.span_bytes, cudaMemcpyHostToDevice, stream) != cudaSuccess) {
            (void)cudaGetLastError();
            return JPEGGPU_INTERNAL_ERROR;
        }
    } else {
        for (int i = 0; i < n; ++i) {
            const jpeggpu_b200_batch::Image& im = *B.images[i];
            if (!im.ok || desc[i].ctas[kBkDestuff] == 0) continue;
            const ScanPlan& sp = im.plan.scans[0];
            if (cudaMemcpyAsync(base + im.layout.off_bytes, items[i].data + sp.begin, sp.geom.scan_avail,
                                cudaMemcpyHostToDevice, stream) != cudaSuccess) {
                (void)cudaGetLastError();
                return JPEGGPU_INTERNAL_ERROR;
            }
        }
    }
    // the staging block may be rewritten once its copy has executed
    if (B.pinned) {
        int dev = -1;
        bool ok = cudaGetDevice(&dev) == cudaSuccess;
        if (ok && B.have_event && B.event_device != dev) {
            cudaEventDestroy(B.copied);
            B.have_event = false;
        }
        if (ok && !B.have_event) {
            ok = cudaEventCreateWithFlags(&B.copied, cudaEventDisableTiming | cudaEventBlockingSync) == cudaSuccess;
            B.have_event = ok;
            B.event_device = dev;
        }
        if (ok && cudaEventRecord(B.copied, stream) == cudaSuccess) {
            B.in_flight = true;
        } else {
            (void)cudaGetLastError();
            if (cudaStreamSynchronize(stream) != cudaSuccess) return JPEGGPU_INTERNAL_ERROR;
        }
    }
    if (launch_batch(bl, write_warps, stream) != cudaSuccess) {
        (void)cudaGetLastError();
        return JPEGGPU_INTERNAL_ERROR;
    }
    return first;
}

} // namespace

extern "C" {

enum jpeggpu_status jpeggpu_b200_batch_create(jpeggpu_b200_batch_t* batch)
{
    if (!batch) return JPEGGPU_INVALID_ARGUMENT;
    *batch = new (std::nothrow) jpeggpu_b200_batch();
    return *batch ? JPEGGPU_SUCCESS : JPEGGPU_OUT_OF_HOST_MEMORY;
}

enum jpeggpu_status jpeggpu_b200_batch_destroy(jpeggpu_b200_batch_t batch)
{
    if (!batch) return JPEGGPU_INVALID_ARGUMENT;
    delete batch;
    return JPEGGPU_SUCCESS;
}

enum jpeggpu_status jpeggpu_b200_batch_set_host_arena(jpeggpu_b200_batch_t batch, const void* base, size_t bytes)
{
    if (!batch || (!base && bytes)) return JPEGGPU_INVALID_ARGUMENT;
    // the span copy starts at a 16-byte boundary at or below the first scan
    const uintptr_t lo = (reinterpret_cast<uintptr_t>(base) + 15) & ~(uintptr_t)15;
    batch->host_lo = bytes ? reinterpret_cast<const uint8_t*>(lo) : nullptr;
    batch->host_hi = bytes ? static_cast<const uint8_t*>(base) + bytes : nullptr;
    return JPEGGPU_SUCCESS;
}

enum jpeggpu_status jpeggpu_b200_batch_plan(
    jpeggpu_b200_batch_t batch, struct jpeggpu_b200_batch_item* items, int num_items, size_t* arena_size)
{
    if (!batch || !items || num_items < 0 || num_items > 65535 || !arena_size) return JPEGGPU_INVALID_ARGUMENT;
    try {
        return plan_batch(*batch, items, num_items, arena_size);
    } catch (const std::bad_alloc&) {
        batch->planned = 0;
        return JPEGGPU_OUT_OF_HOST_MEMORY;
    }
}

enum jpeggpu_status jpeggpu_b200_batch_decode(
    jpeggpu_b200_batch_t batch, struct jpeggpu_b200_batch_item* items, int num_items, void* d_arena,
    size_t arena_size, cudaStream_t stream)
{
    if (!batch || !items || !d_arena) return JPEGGPU_INVALID_ARGUMENT;
    return decode_batch(*batch, items, num_items, d_arena, arena_size, stream);
}

enum jpeggpu_status jpeggpu_b200_batch_read_scan_info(
    jpeggpu_b200_batch_t batch, int index, const void* d_arena, cudaStream_t stream,
    struct jpeggpu_b200_scan_info* info)
{
    if (!batch || !d_arena || !info || index < 0 || index >= batch->planned) return JPEGGPU_INVALID_ARGUMENT;
    const jpeggpu_b200_batch::Image& im = *batch->images[index];
    if (!im.ok) return JPEGGPU_INVALID_ARGUMENT;
    static_assert(sizeof(jb2::ScanInfo) == sizeof(jpeggpu_b200_scan_info), "same four words");
    const uint8_t* src = static_cast<const uint8_t*>(d_arena) + im.layout.off_info;
    if (cudaMemcpyAsync(info, src, sizeof(*info), cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
        cudaStreamSynchronize(stream) != cudaSuccess) {
        (void)cudaGetLastError();
        return JPEGGPU_INTERNAL_ERROR;
    }
    return JPEGGPU_SUCCESS;
}

} // extern "C"
