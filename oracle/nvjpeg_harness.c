/*
 * nvjpeg_harness.c — TEST / BASELINE INFRASTRUCTURE, not part of the product.
 *
 * A thin C wrapper around nvJPEG from the local CUDA toolkit so that bench.py --impl reference
 * and tools/compare_configs.py can time it beside the reference's CUDA decoder in the same run
 * (BASELINE.json north_star: "plus nvJPEG from the local toolkit in the same run"), and so that
 * the GPU tests have a third, independent decoder to cross-check pixels against (different IDCT,
 * so max-abs <= 2, never bit-exact).
 *
 * Protocol: the decoupled API with NVJPEG_OUTPUT_UNCHANGED (native subsampled planes, no colour
 * conversion), one iteration = stream parse + host phase + transfer + device phase — the
 * sequence the reference's own nvJPEG benchmark uses (benchmark/benchmark_nvjpeg.hpp:75-185).
 * The backend is a parameter (the reference uses GPU_HYBRID = 2).
 *
 * Built by oracle/Makefile into oracle/_ref/libnvjpeg_harness.so (links libnvjpeg, libcudart).
 */
#include <cuda_runtime.h>
#include <nvjpeg.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

struct nvjh {
    nvjpegHandle_t handle;
    nvjpegJpegDecoder_t decoder;
    nvjpegJpegState_t state;
    nvjpegBufferDevice_t dev_buf;
    nvjpegBufferPinned_t pin_buf;
    nvjpegJpegStream_t jstream;
    nvjpegDecodeParams_t params;
    int backend;
};

#define TRY(x)                         \
    do {                               \
        nvjpegStatus_t s_ = (x);       \
        if (s_ != NVJPEG_STATUS_SUCCESS) { \
            *status = (int)s_;         \
            goto fail;                 \
        }                              \
    } while (0)

/* Returns NULL and the nvJPEG status in *status if the backend cannot be created. */
void* nvjh_create(int backend, int* status)
{
    struct nvjh* h = (struct nvjh*)calloc(1, sizeof(struct nvjh));
    int dummy;
    if (!status) status = &dummy;
    *status = 0;
    if (!h) {
        *status = -1;
        return NULL;
    }
    h->backend = backend;
    TRY(nvjpegCreateEx((nvjpegBackend_t)backend, NULL, NULL, 0, &h->handle));
    TRY(nvjpegDecoderCreate(h->handle, (nvjpegBackend_t)backend, &h->decoder));
    TRY(nvjpegDecoderStateCreate(h->handle, h->decoder, &h->state));
    TRY(nvjpegBufferDeviceCreate(h->handle, NULL, &h->dev_buf));
    TRY(nvjpegBufferPinnedCreate(h->handle, NULL, &h->pin_buf));
    TRY(nvjpegJpegStreamCreate(h->handle, &h->jstream));
    TRY(nvjpegDecodeParamsCreate(h->handle, &h->params));
    TRY(nvjpegStateAttachDeviceBuffer(h->state, h->dev_buf));
    TRY(nvjpegStateAttachPinnedBuffer(h->state, h->pin_buf));
    TRY(nvjpegDecodeParamsSetOutputFormat(h->params, NVJPEG_OUTPUT_UNCHANGED));
    return h;
fail:
    /* partial teardown is not worth the code in a test harness: leak the few host objects */
    free(h);
    return NULL;
}

void nvjh_destroy(void* p)
{
    struct nvjh* h = (struct nvjh*)p;
    if (!h) return;
    nvjpegDecodeParamsDestroy(h->params);
    nvjpegJpegStreamDestroy(h->jstream);
    nvjpegBufferPinnedDestroy(h->pin_buf);
    nvjpegBufferDeviceDestroy(h->dev_buf);
    nvjpegJpegStateDestroy(h->state);
    nvjpegDecoderDestroy(h->decoder);
    nvjpegDestroy(h->handle);
    free(h);
}

/* Native plane sizes as nvJPEG reports them. */
int nvjh_info(void* p, const uint8_t* data, size_t size, int* channels, int widths[4], int heights[4])
{
    struct nvjh* h = (struct nvjh*)p;
    nvjpegChromaSubsampling_t ss;
    int w[NVJPEG_MAX_COMPONENT], hh[NVJPEG_MAX_COMPONENT];
    nvjpegStatus_t s = nvjpegGetImageInfo(h->handle, data, size, channels, &ss, w, hh);
    if (s != NVJPEG_STATUS_SUCCESS) return (int)s;
    for (int c = 0; c < 4; ++c) {
        widths[c]  = w[c];
        heights[c] = hh[c];
    }
    return 0;
}

/* One decode: parse + host phase + H2D + device phase, all on `stream`; sync != 0 adds a
 * cudaStreamSynchronize.  d_planes[c] are device pointers with row pitch[c] bytes. */
int nvjh_decode(
    void* p, const uint8_t* data, size_t size, uint8_t* d_planes[4], const int pitch[4],
    cudaStream_t stream, int sync)
{
    struct nvjh* h = (struct nvjh*)p;
    nvjpegImage_t img;
    nvjpegStatus_t s;
    memset(&img, 0, sizeof(img));
    for (int c = 0; c < 4; ++c) {
        img.channel[c] = d_planes[c];
        img.pitch[c]   = (size_t)pitch[c];
    }
    s = nvjpegJpegStreamParse(h->handle, data, size, 0, 0, h->jstream);
    if (s != NVJPEG_STATUS_SUCCESS) return (int)s;
    s = nvjpegDecodeJpegHost(h->handle, h->decoder, h->state, h->params, h->jstream);
    if (s != NVJPEG_STATUS_SUCCESS) return (int)s;
    s = nvjpegDecodeJpegTransferToDevice(h->handle, h->decoder, h->state, h->jstream, stream);
    if (s != NVJPEG_STATUS_SUCCESS) return (int)s;
    s = nvjpegDecodeJpegDevice(h->handle, h->decoder, h->state, &img, stream);
    if (s != NVJPEG_STATUS_SUCCESS) return (int)s;
    if (sync && cudaStreamSynchronize(stream) != cudaSuccess) return -2;
    return 0;
}
