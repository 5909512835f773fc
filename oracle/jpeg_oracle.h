/*
 * jpeg_oracle.h — CPU restatement of the jpeggpu decode path.  TEST INFRASTRUCTURE ONLY.
 *
 * Nothing in the product (jpeggpu_b200/, include/) may link or call this.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it, as the
 * checker or as the timed CPU baseline, never as the thing shipped.
 *
 * Each function cites the reference file:line (relative to /root/reference) it restates.
 * Parity status: PINNED on the GPU box against the reference's own CUDA decoder
 * (oracle/_ref/libjpeggpu_ref.so, built from the reference sources by oracle/Makefile) in
 * tests/test_parity_ref.py, and against libjpeg-turbo (Pillow) within max-abs 2 on CPU.
 */
#ifndef JPEG_ORACLE_H_
#define JPEG_ORACLE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAX_COMP 4
#define ORC_MAX_SCANS 4

/* Same numbering as enum jpeggpu_status (include/jpeggpu/jpeggpu.h:38-52). */
enum orc_status {
    ORC_SUCCESS = 0,
    ORC_INVALID_ARGUMENT,
    ORC_INVALID_JPEG,
    ORC_INTERNAL_ERROR,
    ORC_NOT_SUPPORTED,
    ORC_OUT_OF_HOST_MEMORY,
    ORC_INCOMPLETE_BITSTREAM
};

struct orc_scan_info {
    int num_components;           /* components in this scan */
    int comp_idx[ORC_MAX_COMP];   /* frame component index of each scan component */
    int dc_tbl[ORC_MAX_COMP];
    int ac_tbl[ORC_MAX_COMP];
    int begin, end;               /* byte range of the entropy-coded data in the file */
    int dus_per_mcu;
    int mcus_x, mcus_y;
    int data_size_x[ORC_MAX_COMP]; /* per scan component, rounded up to the MCU */
    int data_size_y[ORC_MAX_COMP];
    int num_segments;
    int restart_interval;         /* in force at this SOS (T.81 B.2.4.4: DRI may be redefined between scans) */
};

struct orc_info {
    int width, height, num_components;
    int comp_id[ORC_MAX_COMP];
    int ssx[ORC_MAX_COMP], ssy[ORC_MAX_COMP], qidx[ORC_MAX_COMP];
    int size_x[ORC_MAX_COMP], size_y[ORC_MAX_COMP]; /* output plane sizes */
    int ss_max_x, ss_max_y;
    int restart_interval;
    int num_scans;
    struct orc_scan_info scans[ORC_MAX_SCANS];
    uint8_t qtables[4][64];       /* natural order (reader.cpp:534-545) */
};

/* Marker parse only.  Restates src/reader.cpp:596-672 and the read_* helpers. */
int orc_parse(const uint8_t* data, size_t size, struct orc_info* info);

/* Destuff one scan the way src/decode_destuff.cu:37-113 defines it.
 * in/n: the stuffed entropy-coded bytes of one scan (scan.begin..scan.end).
 * subseq_bytes: subsequence size the segments are padded to (reference: 128).
 * out: buffer of at least orc_destuff_bound(n, subseq_bytes) bytes, zero padded per segment.
 * seg_off/seg_len: per segment, byte offset in out and number of data bytes (capacity max_segs).
 * Returns the number of segments, or -1 if max_segs is too small. *out_size = bytes used. */
size_t orc_destuff_bound(size_t n, int subseq_bytes);
int orc_destuff(const uint8_t* in, size_t n, int subseq_bytes, uint8_t* out, size_t* out_size,
                size_t* seg_off, size_t* seg_len, int max_segs);

/* 8x8 dequantise + IDCT + level shift + clamp, literally src/idct.cu:44-95,:172-221.
 * coef: 64 int16 natural order; q: 64 uint8 natural order; out: 8 rows of 8 bytes at `stride`.
 * ref_signed_q != 0 reproduces the reference's `int8_t qval` cast (src/idct.cu:179, quirk Q3). */
void orc_idct_block(const int16_t* coef, const uint8_t* q, uint8_t* out, int stride,
                    int ref_signed_q);

/* Full decode.
 * stream_coef[s]  (optional, may be NULL): per scan, DU-after-DU coded order, 64 natural-order
 *                 int16 per DU, DC already un-differenced (== reference d_scan_out after decode_dc).
 * planes[c]       (optional): per frame component planar-raster int16 coefficients with row
 *                 length plane_stride[c] == scan data_size_x (reference d_image_qdct, §9.2).
 * pixels[c]       (optional): uint8 output planes with pitch[c].
 * flags bit0: ref_signed_q.  Returns orc_status. */
int orc_decode(const uint8_t* data, size_t size, const struct orc_info* info,
               int16_t* const* stream_coef, int16_t* const* planes, uint8_t* const* pixels,
               const int* pitch, int flags);

/* Number of int16 the per-scan stream-order buffer / per-component plane must hold. */
size_t orc_stream_coef_count(const struct orc_info* info, int scan);
size_t orc_plane_coef_count(const struct orc_info* info, int comp, int* stride_out);

/* Planar YCbCr / grey -> interleaved RGB as the reference's callers do it on the CPU: restates
 * conv_to_rgbi, util/util.h:33-107 (nearest-neighbour chroma replication; float matrix
 * r = y + 1.402 (cr-128), g = y - .344136 (cb-128) - .714136 (cr-128), b = y + 1.772 (cb-128);
 * roundf; clamp).  Planes are dense (row length sizes_x[c]) like the reference's; out holds
 * 3 * sizes_x[0] * sizes_y[0] bytes.  ssx/ssy: frame sampling factors.  Returns 0, or 1 where the
 * reference returns EXIT_FAILURE (component count not 1/3, subsampled luma, unequal chroma).
 * Pinned against the reference's own function (oracle/_ref/libref_util.so) in
 * tests/test_oracle.py; compiled with -ffp-contract=off so no FMA contraction can occur. */
int orc_conv_to_rgbi(const int* sizes_x, const int* sizes_y, int num_components, const int* ssx,
                     const int* ssy, const uint8_t* plane0, const uint8_t* plane1,
                     const uint8_t* plane2, uint8_t* out);

/* Statistics helper used by tests: count Huffman symbols of the whole file (0 on error). */
long orc_count_symbols(const uint8_t* data, size_t size);

#ifdef __cplusplus
}
#endif
#endif
