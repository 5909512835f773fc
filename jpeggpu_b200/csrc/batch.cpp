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
