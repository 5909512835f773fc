/*
 * jpeg_oracle.c — sequential CPU restatement of the jpeggpu decode path.
 * TEST INFRASTRUCTURE ONLY (see jpeg_oracle.h).  Plain C99, no dependencies.
 *
 * Deliberate divergences from the reference (documented in DESIGN.md §quirks):
 *  - Huffman tables persist across scans (intended T.81 semantics; the reference loses them,
 *    SURVEY Q1 / src/reader.cpp:434-441).
 *  - Non-interleaved scans always use 1 data unit per MCU (T.81 A.2.3; reference Q2).
 *  - Writes are bounded by the true data-unit count of the scan (reference Q5).
 */
#include "jpeg_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* zig-zag index -> natural (row-major) index; T.81 Figure A.6, as src/defs.hpp:94-102 */
static const uint8_t k_natural[64] = {
    0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
    41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
    30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

/* ---------------------------------------------------------------- Huffman tables */

struct orc_huff {
    int defined;
    int32_t maxcode[16]; /* largest code of length l+1, -1 if none (reader.cpp:212-223) */
    int32_t valoff[16];  /* valptr - mincode */
    uint8_t huffval[256];
};

/* src/reader.cpp:186-224 (canonical code assignment, T.81 Annex C) */
static void build_huff(struct orc_huff* t, const uint8_t counts[16])
{
    uint16_t huffcode[257];
    int k = 0;
    unsigned code = 0;
    for (int l = 0; l < 16; ++l) {
        for (int i = 0; i < counts[l]; ++i) huffcode[k++] = (uint16_t)code++;
        code <<= 1;
    }
    k = 0;
    for (int l = 0; l < 16; ++l) {
        if (counts[l]) {
            t->valoff[l] = k - huffcode[k];
            k += counts[l];
            t->maxcode[l] = huffcode[k - 1];
        } else {
            t->maxcode[l] = -1;
            t->valoff[l] = 0;
        }
    }
    t->defined = 1;
}

/* ---------------------------------------------------------------- parser */

struct parser {
    const uint8_t* p;
    const uint8_t* begin;
    const uint8_t* end;
    int found_sof;
    int q_defined[4];
    struct orc_huff huff[ORC_MAX_SCANS][8]; /* snapshot per scan; idx = th*2 + class */
    struct orc_huff cur[8];
};

static int remaining(const struct parser* ps) { return (int)(ps->end - ps->p); }
static unsigned rd8(struct parser* ps) { return *ps->p++; }
static unsigned rd16(struct parser* ps)
{
    unsigned h = rd8(ps);
    return (h << 8) | rd8(ps);
}
static int ceil_div(int a, int b) { return (a + b - 1) / b; }

/* src/reader.cpp:81-184 */
static int parse_sof(struct parser* ps, struct orc_info* info)
{
    if (remaining(ps) < 2) return ORC_INVALID_JPEG;
    int length = (int)rd16(ps);
    if (length < 2) return ORC_INVALID_JPEG;
    if (remaining(ps) < length - 2) return ORC_INCOMPLETE_BITSTREAM;
    if (length < 8) return ORC_INVALID_JPEG;
    if (rd8(ps) != 8) return ORC_NOT_SUPPORTED;
    int lines = (int)rd16(ps), samples = (int)rd16(ps);
    if (lines == 0 || samples == 0) return ORC_INVALID_JPEG;
    info->width = samples;
    info->height = lines;
    int nc = (int)rd8(ps);
    if (nc == 0) return ORC_INVALID_JPEG;
    if (nc > ORC_MAX_COMP) return ORC_NOT_SUPPORTED;
    info->num_components = nc;
    if (remaining(ps) < 3 * nc) return ORC_INCOMPLETE_BITSTREAM;
    info->ss_max_x = info->ss_max_y = 0;
    for (int c = 0; c < nc; ++c) {
        info->comp_id[c] = (int)rd8(ps);
        int sf = (int)rd8(ps);
        int sx = sf >> 4, sy = sf & 15;
        if (sx < 1 || sx > 4 || sy < 1 || sy > 4) return ORC_INVALID_JPEG;
        if (nc == 1) sx = sy = 1; /* reader.cpp:147-157 */
        info->ssx[c] = sx;
        info->ssy[c] = sy;
        int qi = (int)rd8(ps);
        if (qi > 3) return ORC_INVALID_JPEG;
        info->qidx[c] = qi;
        if (sx > info->ss_max_x) info->ss_max_x = sx;
        if (sy > info->ss_max_y) info->ss_max_y = sy;
    }
    for (int c = 0; c < nc; ++c) { /* T.81 A.1.1, reader.cpp:175-181 */
        info->size_x[c] = ceil_div(info->width * info->ssx[c], info->ss_max_x);
        info->size_y[c] = ceil_div(info->height * info->ssy[c], info->ss_max_y);
    }
    /* skip any excess bytes the length field declares */
    int used = 8 + 3 * nc;
    if (length > used) ps->p += length - used;
    return ORC_SUCCESS;
}

/* src/reader.cpp:226-303 */
static int parse_dht(struct parser* ps)
{
    if (remaining(ps) < 2) return ORC_INVALID_JPEG;
    int length = (int)rd16(ps) - 2;
    if (length < 0 || remaining(ps) < length) return ORC_INVALID_JPEG;
    int rem = length;
    while (rem > 0) {
        int idx = (int)rd8(ps);
        --rem;
        int cls = idx >> 4, th = idx & 15;
        if (cls != 0 && cls != 1) return ORC_INVALID_JPEG;
        if (th > 3) return ORC_NOT_SUPPORTED;
        if (remaining(ps) < 16) return ORC_INVALID_JPEG;
        uint8_t counts[16];
        int count = 0;
        for (int i = 0; i < 16; ++i) {
            counts[i] = (uint8_t)rd8(ps);
            count += counts[i];
        }
        rem -= 16;
        if (remaining(ps) < count) return ORC_INVALID_JPEG;
        if (count > 256) return ORC_INVALID_JPEG;
        struct orc_huff* t = &ps->cur[th * 2 + cls];
        memset(t, 0, sizeof(*t));
        for (int i = 0; i < count; ++i) t->huffval[i] = (uint8_t)rd8(ps);
        rem -= count;
        build_huff(t, counts);
    }
    return ORC_SUCCESS;
}

/* src/reader.cpp:494-549 */
static int parse_dqt(struct parser* ps, struct orc_info* info)
{
    if (remaining(ps) < 2) return ORC_INVALID_JPEG;
    int length = (int)rd16(ps) - 2;
    if (length < 0 || remaining(ps) < length) return ORC_INVALID_JPEG;
    int rem = length;
    while (rem > 0) {
        int b = (int)rd8(ps);
        --rem;
        int prec = b >> 4, id = b & 15;
        if ((prec != 0 && prec != 1) || id > 3) return ORC_INVALID_JPEG;
        if (prec != 0) return ORC_NOT_SUPPORTED;
        if (rem < 64) return ORC_INVALID_JPEG;
        ps->q_defined[id] = 1;
        for (int j = 0; j < 64; ++j) info->qtables[id][k_natural[j]] = (uint8_t)rd8(ps);
        rem -= 64;
    }
    return ORC_SUCCESS;
}

/* src/reader.cpp:551-574 */
static int parse_dri(struct parser* ps, struct orc_info* info)
{
    if (remaining(ps) < 2) return ORC_INVALID_JPEG;
    int length = (int)rd16(ps) - 2;
    if (length < 2 || remaining(ps) < length) return ORC_INVALID_JPEG;
    int rsti = (int)rd16(ps);
    /* the reference refuses a second, different DRI (src/reader.cpp:563-568); T.81 B.2.4.4 lets the
     * interval change between scans, and each scan uses the value in force at its SOS */
    info->restart_interval = rsti;
    ps->p += length - 2;
    return ORC_SUCCESS;
}

/* src/reader.cpp:576-594 */
static int skip_segment(struct parser* ps)
{
    if (remaining(ps) < 2) return ORC_INVALID_JPEG;
    int length = (int)rd16(ps);
    if (length < 2) return ORC_INVALID_JPEG;
    if (remaining(ps) < length - 2) return ORC_INCOMPLETE_BITSTREAM;
    ps->p += length - 2;
    return ORC_SUCCESS;
}

/* src/reader.cpp:305-492 */
static int parse_sos(struct parser* ps, struct orc_info* info)
{
    if (!ps->found_sof) return ORC_INVALID_JPEG;
    if (remaining(ps) < 3) return ORC_INVALID_JPEG;
    int length = (int)rd16(ps);
    if (length < 3) return ORC_INVALID_JPEG;
    int ns = (int)rd8(ps);
    if (ns < 1 || ns > 4) return ORC_INVALID_JPEG;
    if (info->num_scans >= ORC_MAX_SCANS) return ORC_INVALID_JPEG;
    int si = info->num_scans++;
    struct orc_scan_info* sc = &info->scans[si];
    memset(sc, 0, sizeof(*sc));
    sc->num_components = ns;
    int rest = 2 * ns + 3;
    if (length - 3 != rest) return ORC_INVALID_JPEG;
    if (remaining(ps) < rest) return ORC_INCOMPLETE_BITSTREAM;
    int interleaved = ns > 1;
    for (int a = 0; a < ns; ++a) {
        int sel = (int)rd8(ps);
        int tb = (int)rd8(ps);
        int dc = tb >> 4, ac = tb & 15;
        int ci = -1;
        for (int i = 0; i < info->num_components; ++i)
            if (info->comp_id[i] == sel) {
                ci = i;
                break;
            }
        if (ci < 0) return ORC_INVALID_JPEG;
        if (a > 0 && ci <= sc->comp_idx[a - 1]) return ORC_INVALID_JPEG;
        /* a component may be coded by one scan only (reader.cpp:651-661 asserts it) */
        for (int s = 0; s < si; ++s)
            for (int b = 0; b < info->scans[s].num_components; ++b)
                if (info->scans[s].comp_idx[b] == ci) return ORC_INVALID_JPEG;
        if (dc > 3 || ac > 3) return ORC_INVALID_JPEG;
        if (!ps->cur[dc * 2 + 0].defined) return ORC_INVALID_JPEG;
        if (!ps->cur[ac * 2 + 1].defined) return ORC_INVALID_JPEG;
        if (!ps->q_defined[info->qidx[ci]]) return ORC_INVALID_JPEG;
        sc->comp_idx[a] = ci;
        sc->dc_tbl[a] = dc;
        sc->ac_tbl[a] = ac;
        int mx = interleaved ? 8 * info->ssx[ci] : 8;
        int my = interleaved ? 8 * info->ssy[ci] : 8;
        sc->data_size_x[a] = ceil_div(info->size_x[ci], mx) * mx;
        sc->data_size_y[a] = ceil_div(info->size_y[ci], my) * my;
        sc->mcus_x = sc->data_size_x[a] / mx;
        sc->mcus_y = sc->data_size_y[a] / my;
        sc->dus_per_mcu += interleaved ? info->ssx[ci] * info->ssy[ci] : 1;
    }
    if (sc->dus_per_mcu > 10) return ORC_INVALID_JPEG;
    rd8(ps);
    rd8(ps);
    rd8(ps);
    memcpy(ps->huff[si], ps->cur, sizeof(ps->cur));
    sc->restart_interval = info->restart_interval;

    /* walk the entropy-coded data: reader.cpp:443-489 */
    sc->begin = (int)(ps->p - ps->begin);
    for (;;) {
        const uint8_t* ff = (const uint8_t*)memchr(ps->p, 0xff, (size_t)(ps->end - ps->p));
        if (!ff || ff + 1 >= ps->end) return ORC_INVALID_JPEG;
        unsigned m = ff[1];
        ps->p = ff + 2;
        if (m == 0) continue;
        if (m >= 0xd0 && m <= 0xd7) {
            ++sc->num_segments;
            continue;
        }
        ps->p = ff; /* rewind: end of scan */
        ++sc->num_segments;
        break;
    }
    sc->end = (int)(ps->p - ps->begin);
    return ORC_SUCCESS;
}

static int parse_all(struct parser* ps, struct orc_info* info)
{
    memset(info, 0, sizeof(*info));
    if (remaining(ps) < 2) return ORC_INVALID_JPEG;
    if (rd8(ps) != 0xff) return ORC_INVALID_JPEG;
    if (rd8(ps) != 0xd8) return ORC_INVALID_JPEG;
    for (;;) {
        if (remaining(ps) < 2) return ORC_INVALID_JPEG;
        if (rd8(ps) != 0xff) return ORC_INVALID_JPEG;
        unsigned m = rd8(ps);
        int st = ORC_SUCCESS;
        if (m == 0xc0 || m == 0xc1) {
            if (ps->found_sof) return ORC_INVALID_JPEG;
            ps->found_sof = 1;
            st = parse_sof(ps, info);
        } else if ((m >= 0xc2 && m <= 0xcf) && m != 0xc4 && m != 0xc8 && m != 0xcc) {
            return ORC_NOT_SUPPORTED;
        } else if (m == 0xc4) {
            st = parse_dht(ps);
        } else if (m == 0xd9) {
            break;
        } else if (m == 0xda) {
            st = parse_sos(ps, info);
        } else if (m == 0xdb) {
            st = parse_dqt(ps, info);
        } else if (m == 0xdd) {
            st = parse_dri(ps, info);
        } else {
            st = skip_segment(ps);
        }
        if (st != ORC_SUCCESS) return st;
    }
    if (!ps->found_sof) return ORC_INVALID_JPEG;
    int found[ORC_MAX_COMP] = {0, 0, 0, 0};
    for (int s = 0; s < info->num_scans; ++s)
        for (int a = 0; a < info->scans[s].num_components; ++a)
            found[info->scans[s].comp_idx[a]] = 1;
    for (int c = 0; c < info->num_components; ++c)
        if (!found[c]) return ORC_INVALID_JPEG;
    return ORC_SUCCESS;
}

int orc_parse(const uint8_t* data, size_t size, struct orc_info* info)
{
    if (!data || !info) return ORC_INVALID_ARGUMENT;
    struct parser* ps = (struct parser*)calloc(1, sizeof(struct parser));
    if (!ps) return ORC_OUT_OF_HOST_MEMORY;
    ps->p = ps->begin = data;
    ps->end = data + size;
    int st = parse_all(ps, info);
    free(ps);
    return st;
}

/* ---------------------------------------------------------------- destuff */

size_t orc_destuff_bound(size_t n, int subseq_bytes)
{
    /* every segment is at least 2 stuffed bytes apart (its RSTn marker) */
    return n + (n / 2 + 2) * (size_t)subseq_bytes;
}

/* src/decode_destuff.cu:37-44 (is_byte_data), :64-71 (restart marker), :99-106 (scatter),
 * src/reader.cpp:470-483 (segment sizes rounded up to whole subsequences). */
int orc_destuff(const uint8_t* in, size_t n, int subseq_bytes, uint8_t* out, size_t* out_size,
                size_t* seg_off, size_t* seg_len, int max_segs)
{
    int seg = 0;
    size_t off = 0, len = 0;
    unsigned prev = 0;
    if (max_segs < 1) return -1;
    seg_off[0] = 0;
    for (size_t i = 0; i < n; ++i) {
        unsigned b = in[i];
        if (prev == 0xff) {
            if (b == 0x00) {
                out[off + len++] = 0xff;
            } else if (b >= 0xd0 && b <= 0xd7) {
                seg_len[seg] = len;
                size_t padded = (len + subseq_bytes - 1) / subseq_bytes * subseq_bytes;
                memset(out + off + len, 0, padded - len);
                off += padded;
                len = 0;
                if (++seg >= max_segs) return -1;
                seg_off[seg] = off;
            }
            /* any other byte after 0xff is not data (fill bytes / stray markers) */
        } else if (b != 0xff) {
            out[off + len++] = (uint8_t)b;
        }
        prev = b;
    }
    seg_len[seg] = len;
    size_t padded = (len + subseq_bytes - 1) / subseq_bytes * subseq_bytes;
    memset(out + off + len, 0, padded - len);
    off += padded;
    *out_size = off;
    return seg + 1;
}

/* ---------------------------------------------------------------- IDCT */

/* src/idct.cu:44-47 */
static inline int16_t unfixh(int32_t x) { return (int16_t)(((uint32_t)x + 0x8000u) >> 16); }
static inline int32_t unfixo(int32_t x)
{
    uint32_t u = (uint32_t)x + 0x1000u;
    return (int32_t)u >> 13; /* arithmetic shift, as the device does */
}
static inline int32_t mul(int32_t a, int32_t b) { return (int32_t)((uint32_t)a * (uint32_t)b); }
static inline int32_t add(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
static inline int32_t sub(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }

/* src/idct.cu:50-95, wrap-around arithmetic made explicit */
static void idct_vector(int32_t* v)
{
    const int32_t cos_1_4 = 0x5a82, sin_1_8 = 0x30fc, cos_1_8 = 0x7642, osin_1_16 = 0x063e,
                  osin_5_16 = 0x1a9b, ocos_1_16 = 0x1f63, ocos_5_16 = 0x11c7;
    int32_t tmp10 = mul(add(v[0], v[4]), cos_1_4);
    int32_t tmp11 = mul(sub(v[0], v[4]), cos_1_4);
    int32_t tmp12 = sub(mul(v[2], sin_1_8), mul(v[6], cos_1_8));
    int32_t tmp13 = add(mul(v[6], sin_1_8), mul(v[2], cos_1_8));
    int32_t tmp20 = add(tmp10, tmp13), tmp21 = add(tmp11, tmp12);
    int32_t tmp22 = sub(tmp11, tmp12), tmp23 = sub(tmp10, tmp13);
    int32_t tmp30 = unfixo(mul(add(v[3], v[5]), cos_1_4));
    int32_t tmp31 = unfixo(mul(sub(v[3], v[5]), cos_1_4));
    int32_t v1 = (int32_t)((uint32_t)v[1] << 2), v7 = (int32_t)((uint32_t)v[7] << 2);
    int32_t tmp40 = add(v1, tmp30), tmp41 = add(v7, tmp31);
    int32_t tmp42 = sub(v1, tmp30), tmp43 = sub(v7, tmp31);
    int32_t tmp50 = add(mul(tmp40, ocos_1_16), mul(tmp41, osin_1_16));
    int32_t tmp51 = sub(mul(tmp40, osin_1_16), mul(tmp41, ocos_1_16));
    int32_t tmp52 = add(mul(tmp42, ocos_5_16), mul(tmp43, osin_5_16));
    int32_t tmp53 = sub(mul(tmp42, osin_5_16), mul(tmp43, ocos_5_16));
    v[0] = unfixh(add(tmp20, tmp50));
    v[1] = unfixh(add(tmp21, tmp53));
    v[2] = unfixh(add(tmp22, tmp52));
    v[3] = unfixh(add(tmp23, tmp51));
    v[4] = unfixh(sub(tmp23, tmp51));
    v[5] = unfixh(sub(tmp22, tmp52));
    v[6] = unfixh(sub(tmp21, tmp53));
    v[7] = unfixh(sub(tmp20, tmp50));
}

/* src/idct.cu:172-221: dequantise (int16 truncation), column pass, row pass, +128, clamp */
void orc_idct_block(const int16_t* coef, const uint8_t* q, uint8_t* out, int stride,
                    int ref_signed_q)
{
    int16_t blk[64];
    for (int i = 0; i < 64; ++i) {
        int32_t qv = ref_signed_q ? (int32_t)(int8_t)q[i] : (int32_t)q[i];
        blk[i] = (int16_t)((int32_t)coef[i] * qv);
    }
    int32_t v[8];
    for (int x = 0; x < 8; ++x) {
        for (int y = 0; y < 8; ++y) v[y] = blk[y * 8 + x];
        idct_vector(v);
        for (int y = 0; y < 8; ++y) blk[y * 8 + x] = (int16_t)v[y];
    }
    for (int y = 0; y < 8; ++y) {
        for (int x = 0; x < 8; ++x) v[x] = blk[y * 8 + x];
        idct_vector(v);
        for (int x = 0; x < 8; ++x) {
            int16_t s = (int16_t)(v[x] + 128);
            out[y * stride + x] = (uint8_t)(s < 0 ? 0 : (s > 255 ? 255 : s));
        }
    }
}

/* ---------------------------------------------------------------- entropy decode */

struct bitrd {
    const uint8_t* p;
    size_t nbytes;
    size_t bitpos;
};

/* MSB-first peek of 32 bits, zero-filled past the end of the segment
 * (src/decode_huffman_reader.hpp:72-112 zero-fills the same way). */
static uint32_t peek32(const struct bitrd* br)
{
    size_t byte = br->bitpos >> 3;
    unsigned sh = (unsigned)(br->bitpos & 7);
    uint64_t acc = 0;
    for (int i = 0; i < 5; ++i) {
        uint64_t b = (byte + (size_t)i < br->nbytes) ? br->p[byte + (size_t)i] : 0;
        acc = (acc << 8) | b;
    }
    return (uint32_t)((acc << sh) >> 8);
}

/* src/decode_huffman.cu:167-194 (get_category): canonical length search; an invalid code is
 * forced to terminate at 16 bits with the index wrapped into [0,256). */
static unsigned huff_symbol(const struct orc_huff* t, uint32_t data, int* length)
{
    int i;
    int32_t code = 0;
    for (i = 0; i < 16; ++i) {
        code = (int32_t)(data >> (31 - i));
        if (code <= t->maxcode[i] || i == 15) break;
    }
    *length = i + 1;
    return t->huffval[(uint8_t)(t->valoff[i] + code)];
}

/* src/decode_huffman.cu:196-200 (get_value): T.81 F.2.2.1 EXTEND */
static int extend(int nbits, int code)
{
    return code < ((1 << nbits) >> 1) ? code - (1 << nbits) + 1 : code;
}

size_t orc_stream_coef_count(const struct orc_info* info, int scan)
{
    const struct orc_scan_info* sc = &info->scans[scan];
    return (size_t)sc->mcus_x * sc->mcus_y * sc->dus_per_mcu * 64;
}

size_t orc_plane_coef_count(const struct orc_info* info, int comp, int* stride_out)
{
    for (int s = 0; s < info->num_scans; ++s)
        for (int a = 0; a < info->scans[s].num_components; ++a)
            if (info->scans[s].comp_idx[a] == comp) {
                if (stride_out) *stride_out = info->scans[s].data_size_x[a];
                return (size_t)info->scans[s].data_size_x[a] * info->scans[s].data_size_y[a];
            }
    return 0;
}

static long g_symbols; /* statistics only */

/* Decode one scan into stream order (DU after DU, natural order, DC differential).
 * src/decode_huffman.cu:202-259 (symbols), :360-392 (placement and z/c bookkeeping). */
static int decode_scan_stream(const uint8_t* data, const struct orc_info* info,
                              const struct parser* ps, int si, int16_t* out)
{
    const struct orc_scan_info* sc = &info->scans[si];
    const size_t n = (size_t)(sc->end - sc->begin);
    const int K = sc->dus_per_mcu;
    const long total_mcus = (long)sc->mcus_x * sc->mcus_y;
    const int ri = sc->restart_interval;
    /* which scan component each DU slot of the MCU belongs to */
    int slot_comp[10];
    {
        int k = 0;
        for (int a = 0; a < sc->num_components; ++a) {
            int ci = sc->comp_idx[a];
            int cnt = sc->num_components > 1 ? info->ssx[ci] * info->ssy[ci] : 1;
            for (int j = 0; j < cnt; ++j) slot_comp[k++] = a;
        }
    }
    uint8_t* dst = (uint8_t*)malloc(n + 8);
    int max_segs = sc->num_segments + 1;
    size_t* seg_off = (size_t*)malloc(sizeof(size_t) * (size_t)max_segs * 2);
    if (!dst || !seg_off) {
        free(dst);
        free(seg_off);
        return ORC_OUT_OF_HOST_MEMORY;
    }
    size_t* seg_len = seg_off + max_segs;
    size_t used = 0;
    int nseg = orc_destuff(data + sc->begin, n, 1, dst, &used, seg_off, seg_len, max_segs);
    if (nseg < 0) {
        free(dst);
        free(seg_off);
        return ORC_INTERNAL_ERROR;
    }
    memset(out, 0, orc_stream_coef_count(info, si) * sizeof(int16_t));
    for (int k = 0; k < nseg; ++k) {
        long mcu0 = ri ? (long)k * ri : 0;
        long mcu1 = ri ? mcu0 + ri : total_mcus;
        if (mcu1 > total_mcus) mcu1 = total_mcus;
        if (!ri && k > 0) break;
        struct bitrd br = {dst + seg_off[k], seg_len[k], 0};
        for (long m = mcu0; m < mcu1; ++m) {
            for (int c = 0; c < K; ++c) {
                int a = slot_comp[c];
                const struct orc_huff* tdc = &ps->huff[si][sc->dc_tbl[a] * 2 + 0];
                const struct orc_huff* tac = &ps->huff[si][sc->ac_tbl[a] * 2 + 1];
                int16_t* du = out + ((size_t)m * K + c) * 64;
                int z = 0;
                while (z < 64) {
                    uint32_t d = peek32(&br);
                    int len = 0;
                    unsigned sym = huff_symbol(z == 0 ? tdc : tac, d, &len);
                    int run, cat;
                    if (z == 0) {
                        run = 0;
                        cat = (int)sym;
                    } else {
                        run = (int)(sym >> 4);
                        cat = (int)(sym & 15);
                        if (cat == 0) run = (run == 15) ? 15 : 63 - z;
                    }
                    ++g_symbols;
                    int value = 0;
                    if (cat) {
                        uint32_t rest = d << len;
                        value = extend(cat, (int)(rest >> (32 - cat)));
                        len += cat;
                    }
                    br.bitpos += (size_t)len;
                    z += run;
                    if (value != 0 && z < 64) du[k_natural[z]] = (int16_t)value;
                    ++z;
                }
            }
        }
    }
    free(dst);
    free(seg_off);
    return ORC_SUCCESS;
}

/* src/decode_dc.cu:42-84, :88-169: inclusive sum of DC differences per component in coded
 * order, restarting every restart_interval MCUs.  int16 wrap-around like the device scan. */
static void undo_dc(const struct orc_info* info, int si, int16_t* stream)
{
    const struct orc_scan_info* sc = &info->scans[si];
    const int K = sc->dus_per_mcu;
    const long total_mcus = (long)sc->mcus_x * sc->mcus_y;
    const int ri = sc->restart_interval;
    int off = 0;
    for (int a = 0; a < sc->num_components; ++a) {
        int ci = sc->comp_idx[a];
        int cnt = sc->num_components > 1 ? info->ssx[ci] * info->ssy[ci] : 1;
        int16_t pred = 0;
        for (long m = 0; m < total_mcus; ++m) {
            if (ri && m % ri == 0) pred = 0;
            for (int j = 0; j < cnt; ++j) {
                int16_t* du = stream + ((size_t)m * K + off + j) * 64;
                pred = (int16_t)(pred + du[0]);
                du[0] = pred;
            }
        }
        off += cnt;
    }
}

/* src/decode_transpose.cu:66-131: stream order -> planar raster of 8x8 tiles */
static void transpose_scan(const struct orc_info* info, int si, const int16_t* stream,
                           int16_t* const* planes)
{
    const struct orc_scan_info* sc = &info->scans[si];
    const int K = sc->dus_per_mcu;
    const long total_mcus = (long)sc->mcus_x * sc->mcus_y;
    int off = 0;
    for (int a = 0; a < sc->num_components; ++a) {
        int ci = sc->comp_idx[a];
        int sx = sc->num_components > 1 ? info->ssx[ci] : 1;
        int sy = sc->num_components > 1 ? info->ssy[ci] : 1;
        int stride = sc->data_size_x[a];
        for (long m = 0; m < total_mcus; ++m) {
            int mx = (int)(m % sc->mcus_x), my = (int)(m / sc->mcus_x);
            for (int j = 0; j < sx * sy; ++j) {
                int bx = mx * sx + j % sx, by = my * sy + j / sx;
                const int16_t* du = stream + ((size_t)m * K + off + j) * 64;
                for (int r = 0; r < 8; ++r)
                    memcpy(planes[ci] + (size_t)(by * 8 + r) * stride + bx * 8, du + r * 8, 16);
            }
        }
        off += sx * sy;
    }
}

int orc_decode(const uint8_t* data, size_t size, const struct orc_info* info_in,
               int16_t* const* stream_coef, int16_t* const* planes_out, uint8_t* const* pixels,
               const int* pitch, int flags)
{
    if (!data || !info_in) return ORC_INVALID_ARGUMENT;
    struct parser* ps = (struct parser*)calloc(1, sizeof(struct parser));
    if (!ps) return ORC_OUT_OF_HOST_MEMORY;
    ps->p = ps->begin = data;
    ps->end = data + size;
    struct orc_info info;
    int st = parse_all(ps, &info);
    if (st != ORC_SUCCESS) {
        free(ps);
        return st;
    }
    int16_t* planes[ORC_MAX_COMP] = {0, 0, 0, 0};
    int strides[ORC_MAX_COMP] = {0, 0, 0, 0};
    for (int c = 0; c < info.num_components; ++c) {
        size_t cnt = orc_plane_coef_count(&info, c, &strides[c]);
        planes[c] = (planes_out && planes_out[c]) ? planes_out[c]
                                                  : (int16_t*)malloc(cnt * sizeof(int16_t));
        if (!planes[c]) st = ORC_OUT_OF_HOST_MEMORY;
    }
    for (int s = 0; s < info.num_scans && st == ORC_SUCCESS; ++s) {
        int16_t* stream = (stream_coef && stream_coef[s])
                              ? stream_coef[s]
                              : (int16_t*)malloc(orc_stream_coef_count(&info, s) * sizeof(int16_t));
        if (!stream) {
            st = ORC_OUT_OF_HOST_MEMORY;
            break;
        }
        st = decode_scan_stream(data, &info, ps, s, stream);
        if (st == ORC_SUCCESS) {
            undo_dc(&info, s, stream);
            transpose_scan(&info, s, stream, planes);
        }
        if (!(stream_coef && stream_coef[s])) free(stream);
    }
    if (st == ORC_SUCCESS && pixels) {
        for (int c = 0; c < info.num_components; ++c) {
            if (!pixels[c]) continue;
            int stride = strides[c];
            size_t cnt = orc_plane_coef_count(&info, c, 0);
            int rows = (int)(cnt / (size_t)stride);
            const uint8_t* q = info.qtables[info.qidx[c]];
            for (int by = 0; by < rows / 8; ++by)
                for (int bx = 0; bx < stride / 8; ++bx) {
                    int16_t blk[64];
                    uint8_t px[64];
                    for (int r = 0; r < 8; ++r)
                        memcpy(blk + r * 8, planes[c] + (size_t)(by * 8 + r) * stride + bx * 8, 16);
                    orc_idct_block(blk, q, px, 8, flags & 1);
                    for (int r = 0; r < 8; ++r) {
                        int y = by * 8 + r;
                        if (y >= info.size_y[c]) break;
                        for (int x = 0; x < 8; ++x) {
                            int xx = bx * 8 + x;
                            if (xx < info.size_x[c]) pixels[c][(size_t)y * pitch[c] + xx] = px[r * 8 + x];
                        }
                    }
                }
        }
    }
    for (int c = 0; c < info.num_components; ++c)
        if (!(planes_out && planes_out[c])) free(planes[c]);
    free(ps);
    return st;
}

long orc_count_symbols(const uint8_t* data, size_t size)
{
    struct orc_info info;
    if (orc_parse(data, size, &info) != ORC_SUCCESS) return 0;
    g_symbols = 0;
    if (orc_decode(data, size, &info, 0, 0, 0, 0, 0) != ORC_SUCCESS) return 0;
    return g_symbols;
}

/* util/util.h:33-107.  Written per output pixel instead of per chroma sample: every luma pixel
 * (X, Y) takes the chroma sample (X / ssx, Y / ssy), which is what the reference's nested
 * replication loops visit; the float expression keeps the reference's operation order. */
static uint8_t rgb_sample(float v)
{
    float r = roundf(v); /* util/util.h:97-99: round half away from zero, then clamp */
    if (r < 0.f) r = 0.f;
    if (r > 255.f) r = 255.f;
    return (uint8_t)r;
}

int orc_conv_to_rgbi(const int* sizes_x, const int* sizes_y, int num_components, const int* ssx,
                     const int* ssy, const uint8_t* plane0, const uint8_t* plane1,
                     const uint8_t* plane2, uint8_t* out)
{
    if (num_components != 1 && num_components != 3) return 1; /* :43-46 */
    const int w = sizes_x[0], h = sizes_y[0];
    if (num_components == 1) { /* :48-60 */
        for (size_t i = 0; i < (size_t)w * (size_t)h; ++i)
            out[3 * i] = out[3 * i + 1] = out[3 * i + 2] = plane0[i];
        return 0;
    }
    /* :62-71 */
    if (ssx[0] == 1 && ssy[0] == 1 && (ssx[1] != 1 || ssy[1] != 1) && (ssx[2] != 1 || ssy[2] != 1))
        return 1;
    if (ssx[1] != ssx[2] || ssy[1] != ssy[2]) return 1;
    const int fx = ssx[0], fy = ssy[0]; /* :73-74: luma factors are taken as the replication */
    const int cw = sizes_x[1];          /* :79: chroma row length */
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const size_t ci = (size_t)(y / fy) * (size_t)cw + (size_t)(x / fx);
            const uint8_t cb = plane1[ci], cr = plane2[ci];
            const uint8_t cy = plane0[(size_t)y * (size_t)w + (size_t)x];
            /* :91-95 */
            const float r = cy + 1.402f * (cr - 128.f);
            const float g = cy - .344136f * (cb - 128.f) - .714136f * (cr - 128.f);
            const float b = cy + 1.772f * (cb - 128.f);
            uint8_t* o = out + 3 * ((size_t)y * (size_t)w + (size_t)x);
            o[0] = rgb_sample(r);
            o[1] = rgb_sample(g);
            o[2] = rgb_sample(b);
        }
    return 0;
}
