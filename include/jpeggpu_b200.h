/*
 * jpeggpu_b200.h — additive entry points of libjpeggpu_b200.so (C99).  None of these exist in
 * the reference; the eight jpeggpu_* symbols + is_css_444 of <jpeggpu/jpeggpu.h> are the drop-in
 * boundary and keep working without them.  They expose what the reference keeps private so that
 * parity can be checked through the C ABI:
 *   - the quantised coefficient planes (reference: d_image_qdct inside its scratch,
 *     src/decoder.cpp:230-246, the bit-exact comparison target of SURVEY §9.2),
 *   - per-scan statistics and the scratch offsets of the intermediate buffers (reference:
 *     d_scan_destuffed / s_info at the offsets of SURVEY §9.3).
 */
#ifndef JPEGGPU_B200_H_
#define JPEGGPU_B200_H_

#include <jpeggpu/jpeggpu.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Size in int16 elements of each component's coefficient plane: width[c] x height[c], both
 * rounded up to whole MCUs of the scan that codes the component (reference scan_component::
 * data_size, src/reader.cpp:395-405). */
enum jpeggpu_status jpeggpu_b200_get_coefficient_size(
    jpeggpu_decoder_t decoder, int width[JPEGGPU_MAX_COMP], int height[JPEGGPU_MAX_COMP]);

/* Like jpeggpu_decoder_decode, but stops after DC un-differencing and writes the quantised
 * coefficients (not dequantised, DC absolute) as planar raster int16 into d_planes[c] (device,
 * 16-byte aligned, width[c]*height[c] elements, row length width[c]): the output of the
 * reference's decode_transpose (src/decode_transpose.cu:136-215). */
enum jpeggpu_status jpeggpu_b200_decode_coefficients(
    jpeggpu_decoder_t decoder,
    int16_t* d_planes[JPEGGPU_MAX_COMP],
    void* d_tmp,
    size_t tmp_size,
    cudaStream_t stream);

/* ---- optional output-side fusion: interleaved RGB ---------------------------
 * The reference's contract ends at unconverted planes; its callers convert on the CPU with
 * conv_to_rgbi (util/util.h:33-107: chroma replication, float BT.601 full-range matrix, roundf,
 * clamp; example/example_tool.c calls it before writing the PNG).  These two calls produce
 * byte-identical output on the device.
 *
 * jpeggpu_b200_decode_rgbi: like jpeggpu_decoder_decode, but the IDCT kernel converts in
 * registers and writes height rows of width RGB triples to d_rgb (device, `pitch` bytes per row,
 * pitch >= 3*width); the planes are never materialised.  Accepts what conv_to_rgbi accepts and
 * the fused kernel can see in one pass: one grey component, or Y + 1x1 Cb + 1x1 Cr coded in one
 * interleaved scan; anything else -> JPEGGPU_NOT_SUPPORTED (decode to planes, then convert). */
enum jpeggpu_status jpeggpu_b200_decode_rgbi(
    jpeggpu_decoder_t decoder,
    uint8_t* d_rgb,
    int pitch,
    void* d_tmp,
    size_t tmp_size,
    cudaStream_t stream);

/* conv_to_rgbi on device planes (as produced by jpeggpu_decoder_decode; info as filled by
 * jpeggpu_decoder_parse_header): sizes_x[0] x sizes_y[0] pixels to d_rgb.  1 or 3 components,
 * chroma factors 1x1 (conv_to_rgbi mis-indexes other chroma factors; refused here), else
 * JPEGGPU_NOT_SUPPORTED.  Enqueues one kernel on `stream`. */
enum jpeggpu_status jpeggpu_b200_convert_rgbi(
    const struct jpeggpu_img* planes,
    const struct jpeggpu_img_info* info,
    uint8_t* d_rgb,
    int pitch,
    cudaStream_t stream);

struct jpeggpu_b200_scan_stats {
    int num_components;
    int num_segments;          /* restart segments */
    int dus_per_mcu;
    int total_mcus;
    unsigned scan_bytes;       /* Sb: stuffed entropy-coded bytes */
    unsigned destuffed_bytes;  /* Sd: data bytes after unstuffing */
    unsigned num_subsequences; /* N */
    unsigned num_data_units;   /* D */
    int subsequence_bytes;
    int kernel_launches;       /* launches jpeggpu_decoder_decode issues for this scan */
    /* byte offsets inside d_tmp of the intermediate buffers (valid after decode + sync) */
    size_t off_control, off_destuffed, off_coefficients, off_sync_records, off_slot_prefix;
    /* written by the destuff kernel: the segment directory (16 bytes per restart segment:
     * first subsequence, destuffed bytes, two reserved words) and the totals of the scan
     * (struct jpeggpu_b200_scan_info) */
    size_t off_segments, off_scan_info;
};

/* What the device found in the entropy-coded data of a scan (the destuff kernel locates the end of
 * the scan and the restart markers itself; reference: the host walk src/reader.cpp:443-489 and
 * src/decode_destuff.cu:212-344).  Lives at off_scan_info of the scratch buffer after a decode. */
struct jpeggpu_b200_scan_info {
    unsigned num_subsequences; /* N */
    unsigned num_segments;     /* restart segments found */
    unsigned scan_end;         /* offset, from the first entropy-coded byte, of the byte that ended the scan */
    unsigned status;           /* JPEGGPU_B200_SCAN_* */
};
/* Copies the scan info of scan `scan` of the last decode out of d_tmp (a device-to-host copy on
 * `stream` followed by a synchronisation of that stream: a diagnostic call, not part of the
 * stream-ordered protocol). */
enum jpeggpu_status jpeggpu_b200_read_scan_info(
    jpeggpu_decoder_t decoder, int scan, const void* d_tmp, cudaStream_t stream,
    struct jpeggpu_b200_scan_info* info);
#define JPEGGPU_B200_SCAN_UNTERMINATED 1u /* the data ran out before a terminating marker (no EOI) */
#define JPEGGPU_B200_SCAN_OVERFLOW 2u     /* more restart segments than the frame allows: damaged */

/* Destuff kernel without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Runs the same destuff + restart-segment
 * indexing kernel as jpeggpu_decoder_decode, but the kernel's scratch (look-back descriptors and
 * tile ticket) lives in a static device array inside the library instead of the caller's d_tmp:
 * the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine) and every GPU
 * that loads the library holds 262 400 bytes for it until the process exits.
 *   d_scan          device, 16-byte aligned: the entropy-coded bytes from the first byte behind the
 *                   SOS header; bytes behind the terminating marker may follow (the kernel finds it)
 *   scan_bytes      1 .. JPEGGPU_B200_DESTUFF_STATIC_MAX_BYTES; more -> JPEGGPU_INVALID_ARGUMENT
 *                   and one line on stderr naming the limit
 *   d_destuffed     device, 16-byte aligned, out: 16 * max_subsequences big-endian 32-bit words
 *                   (destuffed byte d of the scan is byte d ^ 3); every restart segment starts on a
 *                   64-byte subsequence and is zero padded to a whole one
 *   d_segments      device, 16-byte aligned, out: max_segments entries
 *   d_info          device, 16-byte aligned, out: totals; JPEGGPU_B200_SCAN_OVERFLOW in status if
 *                   the scan needed more than max_subsequences / max_segments (what fits is written)
 * Enqueues one memset (of the scratch) and one kernel on `stream`; the outputs are valid when the
 * stream has reached that point.  All calls on a device share the one scratch array: calls whose
 * GPU work can overlap (different streams without ordering between them) race and must be
 * serialised by the caller.  Uses the device that is current in the calling thread. */
#define JPEGGPU_B200_DESTUFF_STATIC_MAX_BYTES (64u << 20)
struct jpeggpu_b200_segment {
    unsigned first_subsequence; /* of the segment, in 64-byte subsequences of d_destuffed */
    unsigned num_bytes;         /* destuffed data bytes of the segment */
    unsigned reserved[2];
};
enum jpeggpu_status jpeggpu_b200_destuff_static(
    const uint8_t* d_scan, size_t scan_bytes, uint32_t* d_destuffed, size_t max_subsequences,
    struct jpeggpu_b200_segment* d_segments, size_t max_segments, struct jpeggpu_b200_scan_info* d_info,
    cudaStream_t stream);

/* Statistics of scan `scan` of the last parsed image; *num_scans (optional) gets the count. */
enum jpeggpu_status jpeggpu_b200_get_scan_stats(
    jpeggpu_decoder_t decoder, int scan, struct jpeggpu_b200_scan_stats* stats, int* num_scans);

/* Bytes jpeggpu_decoder_transfer copies host-to-device for the last parsed image. */
enum jpeggpu_status jpeggpu_b200_get_transfer_size(jpeggpu_decoder_t decoder, size_t* bytes);

/* Who walks the entropy-coded bytes (to find the end of the scan and the restart markers).
 *   3 (default)  the device.  For a file whose first scan holds all frame components and at least
 *                64 KiB follow its header, jpeggpu_decoder_parse_header stops at the scan header
 *                (its cost no longer grows with the file), jpeggpu_decoder_get_buffer_size answers
 *                from upper bounds (the scan is planned as if it filled the rest of the file),
 *                jpeggpu_decoder_transfer uploads everything behind the scan header, and the
 *                destuff kernel finds the end of the scan and the restart segments with a
 *                single-pass look-back.  What only the walk can detect - no terminating marker,
 *                stray markers - is then NOT reported by parse_header; it is recorded in struct
 *                jpeggpu_b200_scan_info (the decode stays memory safe and decodes what is there).
 *                Smaller files, and files with several scans, are walked by parse_header.
 *   0            the host, inside jpeggpu_decoder_parse_header, like the reference
 *                (src/reader.cpp:443-489): every status of the reference at parse time.
 *   1            the host, inside jpeggpu_decoder_transfer behind the upload of the bytes (files
 *                that are one scan without DRI and end with EOI; others as 0); statuses of the
 *                walk are returned by transfer.  Scans of at least 1 MiB are walked by the calling
 *                thread plus up to three helper threads owned by the decoder (started on first
 *                use, asleep otherwise; they read the file only while transfer runs).
 *   2            as 1, on the calling thread alone. */
enum jpeggpu_status jpeggpu_b200_set_deferred_walk(jpeggpu_decoder_t decoder, int on);

/* Per-kernel device times. With profiling on, jpeggpu_decoder_decode records a CUDA event on the
 * caller's stream before every kernel; after the caller has synchronised the stream,
 * get_kernel_times returns the milliseconds and launch count per kernel id of the last decode.
 * Profiling is off by default (timed runs never use it). */
#define JPEGGPU_B200_MAX_KERNELS 16
enum jpeggpu_status jpeggpu_b200_set_profiling(jpeggpu_decoder_t decoder, int on);
enum jpeggpu_status jpeggpu_b200_get_kernel_times(
    jpeggpu_decoder_t decoder, float ms[JPEGGPU_B200_MAX_KERNELS],
    int launches[JPEGGPU_B200_MAX_KERNELS]);
const char* jpeggpu_b200_kernel_name(int id);

/* ---- batches (BASELINE config 5: many independent images on one GPU) -------------------------
 * The reference has no batch entry point; its callers loop over the per-image protocol
 * (benchmark/benchmark_jpeggpu.hpp:69-89).  This helper runs exactly that loop natively: one
 * host thread, one decoder and one caller-owned scratch buffer per caller-owned stream ("lane");
 * lane l decodes items l, l + num_lanes, l + 2 num_lanes, ...  Host parsing of one item overlaps
 * the kernels of the previous ones; nothing synchronises with the device except a lane that runs
 * more than four images ahead.  The call returns when every item has been ENQUEUED; the caller
 * synchronises its streams.  All lanes use the device that is current in the calling thread. */
struct jpeggpu_b200_batch_item {
    const uint8_t* data;          /* in: JPEG bytes (pinned host memory recommended) */
    size_t size;                  /* in */
    struct jpeggpu_img img;       /* in: device planes; pitch[c] >= sizes_x[c] */
    struct jpeggpu_img_info info; /* out */
    enum jpeggpu_status status;   /* out: first failing step's status, or JPEGGPU_SUCCESS */
};

/* Returns JPEGGPU_SUCCESS if every item succeeded, else the status of the first failed item
 * (remaining items are still attempted).  An item whose scratch requirement exceeds
 * tmp_size[lane] fails with JPEGGPU_INTERNAL_ERROR. */
enum jpeggpu_status jpeggpu_b200_decode_batch(
    struct jpeggpu_b200_batch_item* items,
    int num_items,
    const cudaStream_t* streams,
    void* const* d_tmp,
    const size_t* tmp_size,
    int num_lanes);

/* ---- fused batches: several images per kernel launch -------------------------------------------
 * The lane helper above still pays, per image, five kernel launches, two copies and an event, and
 * a 2 MP image fills a fraction of the GPU.  A fused batch decodes K images with ONE launch of each
 * of the five kernels (a 2-D grid: image x CTA; per-image descriptors in device memory), one
 * upload of the descriptor/table block, one memset and - when the files lie close together in one
 * host allocation the caller has named (jpeggpu_b200_batch_set_host_arena), e.g. a pinned arena
 * it filled - ONE copy of the whole span of files (otherwise one copy per file).  The device locates the end of every scan and its restart
 * segments itself (walk mode 3 above), so the host only parses headers.
 *
 *   jpeggpu_b200_batch_create(&batch);
 *   per batch:  jpeggpu_b200_batch_plan(batch, items, n, &arena_size);   host only; fills info/status
 *               (caller provides >= arena_size bytes of device memory, 256-byte aligned, and the
 *                output planes in items[i].img)
 *               jpeggpu_b200_batch_decode(batch, items, n, d_arena, arena_size, stream);
 *   jpeggpu_b200_batch_destroy(batch);
 *
 * Everything is enqueued on the caller's stream; nothing synchronises, except that batch_plan
 * waits until the copies of the PREVIOUS decode of the same batch object have left its pinned
 * staging block (use two or three batch objects round-robin to keep the device busy while the
 * host plans).  Items the fused path cannot take - several scans, or any status other than
 * JPEGGPU_SUCCESS at parse time - keep their status in items[i].status and are skipped by
 * batch_decode (decode them with the per-image calls); batch_plan returns the first such status. */
typedef struct jpeggpu_b200_batch* jpeggpu_b200_batch_t;
enum jpeggpu_status jpeggpu_b200_batch_create(jpeggpu_b200_batch_t* batch);
enum jpeggpu_status jpeggpu_b200_batch_destroy(jpeggpu_b200_batch_t batch);
/* Optional: [base, base + bytes) is ONE host allocation (pinned for speed) that holds the files of
 * the coming batches.  Batches whose files lie close together inside it are then uploaded with a
 * single copy of the span from the first scan to the end of the last file; files outside it, or
 * far apart, are copied one by one.  (Without this promise a span could cross unmapped memory.)
 * bytes == 0 withdraws the promise. */
enum jpeggpu_status jpeggpu_b200_batch_set_host_arena(jpeggpu_b200_batch_t batch, const void* base, size_t bytes);
enum jpeggpu_status jpeggpu_b200_batch_plan(
    jpeggpu_b200_batch_t batch, struct jpeggpu_b200_batch_item* items, int num_items, size_t* arena_size);
enum jpeggpu_status jpeggpu_b200_batch_decode(
    jpeggpu_b200_batch_t batch, struct jpeggpu_b200_batch_item* items, int num_items, void* d_arena,
    size_t arena_size, cudaStream_t stream);
/* After the stream has been synchronised: what the device found in item `index` of the last
 * decode (see jpeggpu_b200_read_scan_info; synchronises `stream`). */
enum jpeggpu_status jpeggpu_b200_batch_read_scan_info(
    jpeggpu_b200_batch_t batch, int index, const void* d_arena, cudaStream_t stream,
    struct jpeggpu_b200_scan_info* info);

/* Host-only: parse every item with `num_threads` threads and fill info / status (lets a caller
 * size its output planes); *max_tmp_size (optional) receives the largest scratch requirement. */
enum jpeggpu_status jpeggpu_b200_parse_batch(
    struct jpeggpu_b200_batch_item* items, int num_items, int num_threads, size_t* max_tmp_size);

#ifdef __cplusplus
}
#endif
#endif
