/*
 * jpeggpu.h — C99 interface of the B200-native baseline-JPEG decoder.
 *
 * This header is the drop-in boundary: it declares exactly the symbols, types and enumerator
 * values that nolmoonen/jpeggpu's public header declares (reference include/jpeggpu/jpeggpu.h:
 * 33-111), so a program written against the reference (example/example_tool.c,
 * benchmark/benchmark_jpeggpu.hpp, test/test.cpp) recompiles and relinks against
 * libjpeggpu_b200.so unchanged.  The comments are ours; the ABI is the reference's.
 *
 * Call protocol (reference example/example_tool.c:101-128):
 *   startup -> [ parse_header -> get_buffer_size -> transfer -> decode ]* -> cleanup
 * Everything transfer/decode enqueue is ordered on the caller's stream; no call synchronises
 * the host with the device.  The library never allocates device memory: the caller owns the
 * scratch buffer (>= get_buffer_size bytes, 256-byte aligned), the output planes and the stream.
 */
#ifndef JPEGGPU_JPEGGPU_H_
#define JPEGGPU_JPEGGPU_H_

#include <cuda_runtime.h>

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* reference jpeggpu.h:33 */
#define JPEGGPU_MAX_COMP 4

/* reference jpeggpu.h:35-36 — opaque decoder handle */
struct jpeggpu_decoder;
typedef struct jpeggpu_decoder* jpeggpu_decoder_t;

/* reference jpeggpu.h:38-52 — values 0..6 in this order */
enum jpeggpu_status {
    JPEGGPU_SUCCESS,
    JPEGGPU_INVALID_ARGUMENT,    /* a NULL or out-of-range argument */
    JPEGGPU_INVALID_JPEG,        /* the stream violates ITU-T T.81 */
    JPEGGPU_INTERNAL_ERROR,      /* CUDA failure or scratch buffer too small */
    JPEGGPU_NOT_SUPPORTED,       /* valid JPEG outside baseline 8-bit Huffman */
    JPEGGPU_OUT_OF_HOST_MEMORY,
    JPEGGPU_INCOMPLETE_BITSTREAM /* a marker segment is longer than the remaining bytes */
};

/* reference jpeggpu.h:55 (strings: src/jpeggpu.cpp:41-60) */
const char* jpeggpu_get_status_string(enum jpeggpu_status stat);

/* reference jpeggpu.h:59 — after JPEGGPU_SUCCESS the caller owes one jpeggpu_decoder_cleanup */
enum jpeggpu_status jpeggpu_decoder_startup(jpeggpu_decoder_t* decoder);

/* reference jpeggpu.h:62 — printf logging of the marker walk, off by default */
enum jpeggpu_status jpeggpu_set_logging(jpeggpu_decoder_t decoder, int do_logging);

/* reference jpeggpu.h:65-68 — sampling factors from the frame header, each in [1, 4] */
struct jpeggpu_subsampling {
    int x[JPEGGPU_MAX_COMP];
    int y[JPEGGPU_MAX_COMP];
};

/* reference jpeggpu.h:70 (un-prefixed name kept on purpose) */
int is_css_444(struct jpeggpu_subsampling css, int num_components);

/* reference jpeggpu.h:72-79 — plane sizes are ceil(size * ss / ss_max), unused slots are 0 */
struct jpeggpu_img_info {
    int sizes_x[JPEGGPU_MAX_COMP];
    int sizes_y[JPEGGPU_MAX_COMP];
    int num_components;
    struct jpeggpu_subsampling subsampling;
};

/* reference jpeggpu.h:84-85 — host only.  `data` is borrowed until the copies enqueued by
 * jpeggpu_decoder_transfer have executed; pinned memory makes that copy asynchronous. */
enum jpeggpu_status jpeggpu_decoder_parse_header(
    jpeggpu_decoder_t decoder, struct jpeggpu_img_info* img_info, const uint8_t* data, size_t size);

/* reference jpeggpu.h:88 — bytes of device scratch the parsed image needs */
enum jpeggpu_status jpeggpu_decoder_get_buffer_size(jpeggpu_decoder_t decoder, size_t* tmp_size);

/* reference jpeggpu.h:92-93 — enqueue the host-to-device copies into the front of d_tmp */
enum jpeggpu_status jpeggpu_decoder_transfer(
    jpeggpu_decoder_t decoder, void* d_tmp, size_t tmp_size, cudaStream_t stream);

/* reference jpeggpu.h:97-100 — device planes in frame component order, pitch in bytes */
struct jpeggpu_img {
    uint8_t* image[JPEGGPU_MAX_COMP];
    int pitch[JPEGGPU_MAX_COMP];
};

/* reference jpeggpu.h:104-109 — enqueue the decode kernels; d_tmp is the buffer given to transfer */
enum jpeggpu_status jpeggpu_decoder_decode(
    jpeggpu_decoder_t decoder,
    struct jpeggpu_img* img,
    void* d_tmp,
    size_t tmp_size,
    cudaStream_t stream);

/* reference jpeggpu.h:111 */
enum jpeggpu_status jpeggpu_decoder_cleanup(jpeggpu_decoder_t decoder);

#ifdef __cplusplus
}
#endif

#endif /* JPEGGPU_JPEGGPU_H_ */
