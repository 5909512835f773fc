// parser.cpp — host marker parser and scan planner.  See parser.hpp.
#include "parser.hpp"

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <new>
#include <thread>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace jb2 {

void Logger::log(const char* fmt, ...) const
{
    if (!on) return;
    va_list ap;
    va_start(ap, fmt);
    vprintf(fmt, ap);
    va_end(ap);
}

namespace {

const uint8_t kNatural[64] = {JB2_NATURAL_ORDER};

struct Cursor {
    const uint8_t* p;
    const uint8_t* begin;
    const uint8_t* end;
    size_t left() const { return (size_t)(end - p); }
    unsigned u8() { return *p++; }
    unsigned u16()
    {
        unsigned h = *p++;
        return (h << 8) | *p++;
    }
};

int ceil_div(int a, int b) { return (a + b - 1) / b; }

const char* marker_name(unsigned m)
{
    switch (m) {
    case 0xc0: return "SOF0";
    case 0xc1: return "SOF1";
    case 0xc2: return "SOF2";
    case 0xc4: return "DHT";
    case 0xd8: return "SOI";
    case 0xd9: return "EOI";
    case 0xda: return "SOS";
    case 0xdb: return "DQT";
    case 0xdd: return "DRI";
    case 0xfe: return "COM";
    default: break;
    }
    if (m >= 0xe0 && m <= 0xef) return "APPn";
    if (m >= 0xc3 && m <= 0xcf) return "SOFn";
    return "marker";
}

// Frame header. Statuses follow src/reader.cpp:81-184.
jpeggpu_status read_sof(Cursor& cur, ImagePlan& plan, const Logger& log)
{
    if (cur.left() < 2) return JPEGGPU_INVALID_JPEG;
    const int length = (int)cur.u16();
    if (length < 2) return JPEGGPU_INVALID_JPEG;
    if (cur.left() < (size_t)(length - 2)) return JPEGGPU_INCOMPLETE_BITSTREAM;
    if (length < 8) return JPEGGPU_INVALID_JPEG;
    const uint8_t* body_end = cur.p + (length - 2);
    const unsigned precision = cur.u8();
    if (precision != 8) {
        log.log("\tunsupported sample precision %u, only 8 is supported\n", precision);
        return JPEGGPU_NOT_SUPPORTED;
    }
    const int lines = (int)cur.u16();
    const int samples = (int)cur.u16();
    if (lines == 0 || samples == 0) return JPEGGPU_INVALID_JPEG;
    plan.width = samples;
    plan.height = lines;
    const int nc = (int)cur.u8();
    if (nc == 0) return JPEGGPU_INVALID_JPEG;
    if (nc > kMaxComp) return JPEGGPU_NOT_SUPPORTED;
    if (length - 8 < 3 * nc) return JPEGGPU_INCOMPLETE_BITSTREAM;
    plan.num_comp = nc;
    log.log("\tsize_x: %d, size_y: %d, num_components: %d\n", samples, lines, nc);
    plan.ss_max_x = plan.ss_max_y = 0;
    for (int c = 0; c < nc; ++c) {
        FrameComp& fc = plan.comps[c];
        fc.id = (int)cur.u8();
        const unsigned sf = cur.u8();
        int sx = (int)(sf >> 4), sy = (int)(sf & 15);
        if (sx < 1 || sx > 4 || sy < 1 || sy > 4) return JPEGGPU_INVALID_JPEG;
        if (nc == 1) sx = sy = 1; // a lone component is always 1x1 (src/reader.cpp:147-157)
        fc.ssx = sx;
        fc.ssy = sy;
        fc.qidx = (int)cur.u8();
        if (fc.qidx > 3) return JPEGGPU_INVALID_JPEG;
        log.log("\tc_id: %d, ssx: %d, ssy: %d, qi: %d\n", fc.id, sx, sy, fc.qidx);
        if (sx > plan.ss_max_x) plan.ss_max_x = sx;
        if (sy > plan.ss_max_y) plan.ss_max_y = sy;
    }
    for (int c = 0; c < nc; ++c) {
        FrameComp& fc = plan.comps[c];
        fc.size_x = ceil_div(plan.width * fc.ssx, plan.ss_max_x);
        fc.size_y = ceil_div(plan.height * fc.ssy, plan.ss_max_y);
    }
    cur.p = body_end;
    return JPEGGPU_SUCCESS;
}

// Huffman table definitions. Statuses follow src/reader.cpp:226-303.
jpeggpu_status read_dht(Cursor& cur, ParseState& st, const Logger& log)
{
    if (cur.left() < 2) return JPEGGPU_INVALID_JPEG;
    const int length = (int)cur.u16() - 2;
    if (length < 0 || cur.left() < (size_t)length) return JPEGGPU_INVALID_JPEG;
    int rem = length;
    while (rem > 0) {
        const unsigned tc_th = cur.u8();
        --rem;
        const unsigned cls = tc_th >> 4, th = tc_th & 15;
        if (cls > 1) return JPEGGPU_INVALID_JPEG;
        if (th > 3) return JPEGGPU_NOT_SUPPORTED;
        if (rem < 16) return JPEGGPU_INVALID_JPEG;
        log.log("\t%s Huffman table index %u\n", cls == 0 ? "DC" : "AC", th);
        HuffSpec& h = st.huff[th * 2 + cls];
        h = HuffSpec{};
        int count = 0;
        for (int i = 0; i < 16; ++i) {
            h.counts[i] = (uint8_t)cur.u8();
            count += h.counts[i];
        }
        rem -= 16;
        if (count > 256 || rem < count) return JPEGGPU_INVALID_JPEG;
        // the code lengths must fit the code space (Kraft); otherwise no canonical code exists
        unsigned code = 0;
        for (int l = 0; l < 16; ++l) {
            code += h.counts[l];
            if (code > (1u << (l + 1))) return JPEGGPU_INVALID_JPEG;
            code <<= 1;
        }
        for (int i = 0; i < count; ++i) h.vals[i] = (uint8_t)cur.u8();
        rem -= count;
        h.is_dc = cls == 0;
        h.defined = true;
    }
    return JPEGGPU_SUCCESS;
}

// Quantisation tables, stored de-zigzagged. Statuses follow src/reader.cpp:494-549.
jpeggpu_status read_dqt(Cursor& cur, ImagePlan& plan, const Logger& log)
{
    if (cur.left() < 2) return JPEGGPU_INVALID_JPEG;
    const int length = (int)cur.u16() - 2;
    if (length < 0 || cur.left() < (size_t)length) return JPEGGPU_INVALID_JPEG;
    int rem = length;
    while (rem > 0) {
        const unsigned pq_tq = cur.u8();
        --rem;
        const unsigned pq = pq_tq >> 4, tq = pq_tq & 15;
        if (pq > 1 || tq > 3) return JPEGGPU_INVALID_JPEG;
        if (pq != 0) {
            log.log("\t16-bit quantization table is not supported\n");
            return JPEGGPU_NOT_SUPPORTED;
        }
        if (rem < 64) return JPEGGPU_INVALID_JPEG;
        for (int j = 0; j < 64; ++j) plan.qtables[tq][kNatural[j]] = (uint8_t)cur.u8();
        plan.q_defined[tq] = true;
        rem -= 64;
    }
    return JPEGGPU_SUCCESS;
}

// Restart interval. Statuses follow src/reader.cpp:551-574.
jpeggpu_status read_dri(Cursor& cur, ImagePlan& plan, const Logger& log)
{
    if (cur.left() < 2) return JPEGGPU_INVALID_JPEG;
    const int length = (int)cur.u16() - 2;
    if (length < 2 || cur.left() < (size_t)length) return JPEGGPU_INVALID_JPEG;
    const int rsti = (int)cur.u16();
    // The reference refuses a second, different DRI (src/reader.cpp:563-568).  T.81 B.2.4.4 lets
    // the interval change between scans: every scan takes the value in force at its SOS
    // (ScanGeom::restart_interval, read_sos).
    plan.restart_interval = rsti;
    log.log("\trestart_interval: %d\n", rsti);
    cur.p += length - 2;
    return JPEGGPU_SUCCESS;
}

// APPn / COM / anything else with a length. Statuses follow src/reader.cpp:576-594.
jpeggpu_status skip_segment(Cursor& cur, const Logger& log)
{
    if (cur.left() < 2) return JPEGGPU_INVALID_JPEG;
    const int length = (int)cur.u16();
    if (length < 2) return JPEGGPU_INVALID_JPEG;
    if (cur.left() < (size_t)(length - 2)) return JPEGGPU_INCOMPLETE_BITSTREAM;
    log.log("\twarning: skipping this segment\n");
    cur.p += length - 2;
    return JPEGGPU_SUCCESS;
}

// ---- entropy-coded data walk -------------------------------------------------------------------
// Replaces the memchr walk of src/reader.cpp:443-489 (one libc call per 0xff byte: 336 us on
// IMG_6510.JPG).  Here: one SIMD pass.  In a block of plain bytes every 0xff followed by 0x00 is a
// stuffed pair, which only changes a counter (data bytes = bytes - number of 0xff); the scalar
// path runs only for the 0xff bytes that are NOT followed by 0x00: restart markers, fill bytes
// and the marker that ends the scan.
// The classification is the one the destuff kernel applies per byte (src/decode_destuff.cu:37-44,
// :64-71): a byte is data unless it is 0xff or follows 0xff; 0xff 0x00 contributes one data byte
// (at the position of the 0x00); 0xff 0xd0..0xd7 closes a segment; 0xff followed by anything
// else (after optional 0xff fill bytes) ends the scan.

// Scans [p, lim) for the first 0xff that is not followed by 0x00; returns its address (or lim)
// and adds the number of 0xff bytes before it to *ffs.  Reads at most 1 byte past lim.
typedef const uint8_t* (*FindFn)(const uint8_t* p, const uint8_t* lim, uint64_t* ffs);

const uint8_t* find_marker_scalar(const uint8_t* p, const uint8_t* lim, uint64_t* ffs)
{
    uint64_t n = 0;
    for (; p < lim; ++p)
        if (*p == 0xff) {
            if (p[1] != 0) break;
            ++n;
        }
    *ffs += n;
    return p;
}

#if defined(__x86_64__)
const uint8_t* find_marker_sse2(const uint8_t* p, const uint8_t* lim, uint64_t* ffs)
{
    const __m128i vff = _mm_set1_epi8((char)0xff), vz = _mm_setzero_si128();
    uint64_t n = 0;
    while (p + 16 <= lim) {
        const __m128i a = _mm_loadu_si128((const __m128i*)p);
        const __m128i b = _mm_loadu_si128((const __m128i*)(p + 1));
        const unsigned mf = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(a, vff));
        if (mf) {
            const unsigned mz = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(b, vz));
            const unsigned other = mf & ~mz;
            if (other) {
                const unsigned i = (unsigned)__builtin_ctz(other);
                *ffs += n + (uint64_t)__builtin_popcount(mf & ((1u << i) - 1u));
                return p + i;
            }
            n += (uint64_t)__builtin_popcount(mf);
        }
        p += 16;
    }
    *ffs += n;
    return find_marker_scalar(p, lim, ffs);
}

constexpr int kWalkPrefetch = 4096;

__attribute__((target("avx2,popcnt"))) const uint8_t* find_marker_avx2(const uint8_t* p,
                                                                      const uint8_t* lim,
                                                                      uint64_t* ffs)
{
    // branch-free per 64 bytes: a branch on "block contains 0xff" mispredicts on ~1 block in 8
    const __m256i vff = _mm256_set1_epi8((char)0xff), vz = _mm256_setzero_si256();
    uint64_t n = 0;
    while (p + 64 <= lim) {
        _mm_prefetch((const char*)p + kWalkPrefetch, _MM_HINT_T0);
        const __m256i a0 = _mm256_loadu_si256((const __m256i*)p);
        const __m256i a1 = _mm256_loadu_si256((const __m256i*)(p + 32));
        const __m256i b0 = _mm256_loadu_si256((const __m256i*)(p + 1));
        const __m256i b1 = _mm256_loadu_si256((const __m256i*)(p + 33));
        const uint64_t mf = (uint64_t)(unsigned)_mm256_movemask_epi8(_mm256_cmpeq_epi8(a0, vff)) |
                            ((uint64_t)(unsigned)_mm256_movemask_epi8(_mm256_cmpeq_epi8(a1, vff)) << 32);
        const uint64_t mz = (uint64_t)(unsigned)_mm256_movemask_epi8(_mm256_cmpeq_epi8(b0, vz)) |
                            ((uint64_t)(unsigned)_mm256_movemask_epi8(_mm256_cmpeq_epi8(b1, vz)) << 32);
        const uint64_t other = mf & ~mz;
        if (__builtin_expect(other != 0, 0)) {
            const unsigned i = (unsigned)__builtin_ctzll(other);
            *ffs += n + (uint64_t)__builtin_popcountll(mf & ((1ull << i) - 1ull));
            return p + i;
        }
        n += (uint64_t)__builtin_popcountll(mf);
        p += 64;
    }
    *ffs += n;
    return find_marker_sse2(p, lim, ffs);
}

// 128 bytes per iteration with mask registers: two loads, the "followed by 0x00" mask is derived
// from the zero masks of the same two vectors (+ one scalar byte), no overlapping loads.
__attribute__((target("avx512f,avx512bw,popcnt,avx2"))) const uint8_t* find_marker_avx512(
    const uint8_t* p, const uint8_t* lim, uint64_t* ffs)
{
    const __m512i vff = _mm512_set1_epi8((char)0xff);
    uint64_t n = 0;
    while (p + 128 <= lim) { // p[128] is readable: lim <= end - 1
        // The previous image's H2D copy (a DMA read) leaves these lines out of the core's caches;
        // prefetching 4 KB ahead cut the in-flow walk of a 1.4 MB scan from 60 to 42 us on the
        // B200 host.  (Prefetches past the buffer end do not fault.)
        _mm_prefetch((const char*)p + kWalkPrefetch, _MM_HINT_T0);
        _mm_prefetch((const char*)p + kWalkPrefetch + 64, _MM_HINT_T0);
        const __m512i a0 = _mm512_loadu_si512((const void*)p);
        const __m512i a1 = _mm512_loadu_si512((const void*)(p + 64));
        const uint64_t mf0 = _mm512_cmpeq_epi8_mask(a0, vff);
        const uint64_t mf1 = _mm512_cmpeq_epi8_mask(a1, vff);
        const uint64_t z0 = _mm512_testn_epi8_mask(a0, a0);
        const uint64_t z1 = _mm512_testn_epi8_mask(a1, a1);
        const uint64_t zn0 = (z0 >> 1) | (z1 << 63);
        const uint64_t zn1 = (z1 >> 1) | ((uint64_t)(p[128] == 0) << 63);
        const uint64_t o0 = mf0 & ~zn0, o1 = mf1 & ~zn1;
        if (__builtin_expect((o0 | o1) != 0, 0)) {
            if (o0) {
                const unsigned i = (unsigned)__builtin_ctzll(o0);
                *ffs += n + (uint64_t)__builtin_popcountll(mf0 & ((1ull << i) - 1ull));
                return p + i;
            }
            const unsigned i = (unsigned)__builtin_ctzll(o1);
            *ffs += n + (uint64_t)__builtin_popcountll(mf0) +
                    (uint64_t)__builtin_popcountll(mf1 & ((1ull << i) - 1ull));
            return p + 64 + i;
        }
        n += (uint64_t)__builtin_popcountll(mf0) + (uint64_t)__builtin_popcountll(mf1);
        p += 128;
    }
    *ffs += n;
    return find_marker_avx2(p, lim, ffs);
}

FindFn pick_find_marker()
{
    __builtin_cpu_init();
    if (__builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512f")) return find_marker_avx512;
    return __builtin_cpu_supports("avx2") ? find_marker_avx2 : find_marker_sse2;
}
#else
FindFn pick_find_marker() { return find_marker_scalar; }
#endif

FindFn find_marker_fn()
{
    static const FindFn f = pick_find_marker();
    return f;
}

// start_off (a multiple of kDestuffTile): walk from that offset of the scan as if the bytes
// before it held no data at all; the tiles before it are left as placeholders.  The caller adds
// what the skipped bytes contribute (parallel walk below).
jpeggpu_status walk_scan(Cursor& cur, ScanPlan& sp, uint64_t start_off = 0)
{
    const FindFn find_marker = find_marker_fn();
    const uint8_t* const scan_begin = cur.p;
    const uint8_t* p = cur.p + start_off;
    const uint8_t* const end = cur.end;
    uint64_t kept_total = 0;   // data bytes of the scan before p
    uint64_t seg_kept0 = 0;    // kept_total at the start of the open segment
    uint64_t sub_off = 0, src_begin = 0;
    uint64_t tile_next = start_off; // stuffed offset of the next tile boundary to record
    sp.tiles.resize((size_t)(start_off / kDestuffTile));
    auto close_segment = [&]() {
        SegDesc d;
        d.sub_off = (uint32_t)sub_off;
        d.nbytes = (uint32_t)(kept_total - seg_kept0);
        d.kept_before = (uint32_t)seg_kept0;
        d.src_begin = (uint32_t)src_begin;
        sp.segs.push_back(d);
        sub_off += (kept_total - seg_kept0 + kSubBytes - 1) / kSubBytes;
        seg_kept0 = kept_total;
    };
    auto emit_tile = [&]() { // totals over the bytes before the boundary
        TileDesc t;
        t.kept_before = (uint32_t)kept_total;
        t.segs_before = (uint32_t)sp.segs.size();
        sp.tiles.push_back(t);
        tile_next += kDestuffTile;
    };
    for (;;) {
        if ((uint64_t)(p - scan_begin) == tile_next) emit_tile();
        // plain bytes up to the next tile boundary (the scan ends with a marker, so a 0xff that
        // is not stuffed is found before `end - 1` in every valid file)
        const uint8_t* lim = scan_begin + tile_next;
        if (lim > end - 1) lim = end - 1;
        if (p >= lim && lim == end - 1) return JPEGGPU_INVALID_JPEG; // no EOI
        uint64_t ffs = 0;
        const uint8_t* f = find_marker(p, lim, &ffs);
        kept_total += (uint64_t)(f - p) - ffs;
        p = f;
        if (f == lim) continue; // tile boundary (or the tail check above)
        // f: 0xff not followed by 0x00
        const uint8_t* q = f + 1;
        while (q < end && *q == 0xff) ++q; // fill bytes
        if (q >= end) return JPEGGPU_INVALID_JPEG;
        while (tile_next <= (uint64_t)(q - scan_begin)) emit_tile(); // none of f..q-1 is data
        const unsigned m = *q;
        if (m == 0) { // 0xff 0xff 0x00: still one stuffed data byte
            ++kept_total;
            p = q + 1;
            continue;
        }
        if (m >= 0xd0 && m <= 0xd7) {
            close_segment();
            p = q + 1;
            src_begin = (uint64_t)(p - scan_begin);
            continue;
        }
        close_segment();
        cur.p = q - 1; // the 0xff that introduces the next marker
        break;
    }
    sp.begin = (size_t)(scan_begin - cur.begin);
    sp.end = (size_t)(cur.p - cur.begin);
    if (sp.end - sp.begin >= (1ull << 31) || sub_off >= (1ull << 31)) return JPEGGPU_NOT_SUPPORTED;
    sp.geom.scan_bytes = (uint32_t)(sp.end - sp.begin);
    sp.geom.scan_avail = sp.geom.scan_bytes + 2u; // the marker the walk stopped at
    sp.geom.num_subs = (uint32_t)sub_off;
    sp.geom.num_segments = (int)sp.segs.size();
    sp.tiles.resize(((size_t)sp.geom.scan_bytes + kDestuffTile - 1) / kDestuffTile);
    return JPEGGPU_SUCCESS;
}


// ---- parallel walk (see parser.hpp) ------------------------------------------------------------
constexpr size_t kWalkBlock = 64 * 1024; // counting unit handed to a thread
constexpr int kWalkHelpers = 3;

struct WalkJob {
    const uint8_t* scan_begin = nullptr;
    uint32_t num_blocks = 0;          // counting blocks [0, num_blocks) of kWalkBlock bytes
    std::vector<uint32_t> tile_ffs;   // 0xff bytes per destuff tile of the counting region
    std::atomic<uint32_t> next{0}, done{0};
    std::atomic<bool> impure{false};
};

// Pull counting blocks until none is left.  A block is pure when every 0xff in it is followed by
// 0x00 (the byte after the block included): then data bytes = bytes - number of 0xff.
void walk_job_work(WalkJob& job)
{
    const FindFn find_marker = find_marker_fn();
    constexpr uint32_t tiles_per_block = (uint32_t)(kWalkBlock / kDestuffTile);
    for (;;) {
        const uint32_t b = job.next.fetch_add(1, std::memory_order_relaxed);
        if (b >= job.num_blocks) return;
        const uint8_t* t0 = job.scan_begin + (size_t)b * kWalkBlock;
        for (uint32_t i = 0; i < tiles_per_block; ++i, t0 += kDestuffTile) {
            uint64_t ffs = 0;
            if (find_marker(t0, t0 + kDestuffTile, &ffs) != t0 + kDestuffTile)
                job.impure.store(true, std::memory_order_relaxed);
            job.tile_ffs[(size_t)b * tiles_per_block + i] = (uint32_t)ffs;
        }
        job.done.fetch_add(1, std::memory_order_release);
    }
}

} // namespace

struct WalkPool {
    std::vector<std::thread> threads;
    std::mutex m;
    std::condition_variable cv;
    bool stop = false;
    std::shared_ptr<WalkJob> job; // the job of the image being transferred (guarded by m)
    uint64_t generation = 0;
    uint64_t kicked_for = 0; // ImagePlan::generation of the image the job was cut from (0: none)
    uint64_t stitched = 0; // walks completed on the parallel path (tests, diagnostics)

    void worker()
    {
        uint64_t seen = 0;
        for (;;) {
            std::shared_ptr<WalkJob> mine;
            {
                std::unique_lock<std::mutex> lk(m);
                cv.wait(lk, [&] { return stop || generation != seen; });
                if (stop) return;
                seen = generation;
                mine = job;
            }
            // a helper that wakes late finds its (old) job exhausted and never touches file bytes
            if (mine) walk_job_work(*mine);
        }
    }
};

WalkPool* walk_pool_create() { return new (std::nothrow) WalkPool; }

uint64_t walk_pool_stitched(const WalkPool* pool) { return pool ? pool->stitched : 0; }

void walk_pool_destroy(WalkPool* pool)
{
    if (!pool) return;
    {
        std::lock_guard<std::mutex> lk(pool->m);
        pool->stop = true;
    }
    pool->cv.notify_all();
    for (std::thread& t : pool->threads) t.join();
    delete pool;
}

void walk_pool_kick(WalkPool* pool, const ImagePlan& plan)
{
    if (!pool || !plan.walk_pending) return;
    pool->kicked_for = 0;
    const ScanPlan& sp = plan.scans[0];
    const size_t avail = plan.file_size - sp.begin; // the scan runs to the EOI at the end of the file
    if (avail < kWalkParallelMin) return;
    try {
        if (pool->threads.empty())
            for (int i = 0; i < kWalkHelpers; ++i) pool->threads.emplace_back([pool] { pool->worker(); });
        auto job = std::make_shared<WalkJob>();
        job->scan_begin = plan.file + sp.begin;
        job->num_blocks = (uint32_t)((avail - kWalkBlock) / kWalkBlock); // tail: 64..128 KB
        job->tile_ffs.resize((size_t)job->num_blocks * (kWalkBlock / kDestuffTile));
        {
            std::lock_guard<std::mutex> lk(pool->m);
            pool->job = job;
            ++pool->generation;
        }
        pool->cv.notify_all();
        pool->kicked_for = plan.generation;
    } catch (...) { // no threads / no memory: the sequential walk needs neither
        pool->kicked_for = 0;
    }
}

namespace {
// take every block that has not been handed out; wait for the ones in progress
void drain_job(WalkJob& job)
{
    const uint32_t taken = job.next.exchange(job.num_blocks, std::memory_order_relaxed);
    const uint32_t started = taken < job.num_blocks ? taken : job.num_blocks;
    while (job.done.load(std::memory_order_acquire) < started) std::this_thread::yield();
}
} // namespace

void walk_pool_cancel(WalkPool* pool)
{
    if (!pool || !pool->kicked_for || !pool->job) return;
    pool->kicked_for = 0;
    drain_job(*pool->job);
}

namespace {

// Returns false when the parallel path does not apply or its result had to be discarded (the
// caller then walks sequentially from a clean state); true with *status set otherwise.
bool walk_parallel(WalkPool* pool, const ImagePlan& plan, Cursor& cur, ScanPlan& sp,
                   jpeggpu_status* status)
{
    if (!pool || pool->kicked_for == 0 || pool->kicked_for != plan.generation || !pool->job) return false;
    pool->kicked_for = 0;
    const std::shared_ptr<WalkJob> job = pool->job;
    if (job->scan_begin != cur.p) {
        drain_job(*job);
        return false;
    }
    const uint64_t counted = (uint64_t)job->num_blocks * kWalkBlock;
    // tail first (only this thread can do it), then help with the counting blocks
    sp.segs.clear();
    sp.tiles.clear();
    const jpeggpu_status st = walk_scan(cur, sp, counted);
    walk_job_work(*job);
    while (job->done.load(std::memory_order_acquire) != job->num_blocks) {
#if defined(__x86_64__)
        _mm_pause();
#endif
    }
    if (st != JPEGGPU_SUCCESS || job->impure.load(std::memory_order_relaxed) || sp.segs.size() != 1)
        return false; // let the sequential walk decide (it also produces the right status)
    // stitch: tile totals of the counting region, then shift the tail by what that region holds
    uint64_t kept = 0;
    const size_t counted_tiles = (size_t)(counted / kDestuffTile);
    for (size_t i = 0; i < counted_tiles && i < sp.tiles.size(); ++i) {
        sp.tiles[i].kept_before = (uint32_t)kept;
        sp.tiles[i].segs_before = 0;
        kept += kDestuffTile - job->tile_ffs[i];
    }
    for (size_t i = counted_tiles; i < sp.tiles.size(); ++i) {
        if (sp.tiles[i].segs_before != 0) return false;
        sp.tiles[i].kept_before += (uint32_t)kept;
    }
    SegDesc& seg = sp.segs[0];
    const uint64_t nbytes = (uint64_t)seg.nbytes + kept;
    const uint64_t subs = (nbytes + kSubBytes - 1) / kSubBytes;
    if (nbytes >= (1ull << 31) || subs >= (1ull << 31)) return false;
    seg.nbytes = (uint32_t)nbytes;
    seg.src_begin = 0;
    sp.geom.num_subs = (uint32_t)subs;
    ++pool->stitched;
    *status = JPEGGPU_SUCCESS;
    return true;
}

// Scan header + geometry. Statuses follow src/reader.cpp:305-441.
jpeggpu_status read_sos(Cursor& cur, ImagePlan& plan, ParseState& st, const Logger& log,
                        int walk)
{
    if (!st.found_sof) return JPEGGPU_INVALID_JPEG;
    if (cur.left() < 3) return JPEGGPU_INVALID_JPEG;
    const int length = (int)cur.u16();
    if (length < 3) return JPEGGPU_INVALID_JPEG;
    const int ns = (int)cur.u8();
    if (ns < 1 || ns > 4) return JPEGGPU_INVALID_JPEG;
    if (plan.num_scans >= kMaxScans) return JPEGGPU_INVALID_JPEG;
    const int rest = 2 * ns + 3;
    if (length - 3 != rest) return JPEGGPU_INVALID_JPEG;
    if (cur.left() < (size_t)rest) return JPEGGPU_INCOMPLETE_BITSTREAM;

    ScanPlan& sp = plan.scans[plan.num_scans];
    sp.segs.clear(); // keep the capacity of the directories across images
    sp.tiles.clear();
    sp.begin = sp.end = 0;
    for (int i = 0; i < kMaxTables; ++i) sp.tables[i] = HuffSpec{};
    memset(sp.qidx, 0, sizeof(sp.qidx));
    ScanGeom& g = sp.geom;
    memset(&g, 0, sizeof(g));
    g.num_comp = ns;
    const bool interleaved = ns > 1;
    int table_of[kMaxTables]; // global table index -> slot in this scan, -1 if unused
    for (int i = 0; i < kMaxTables; ++i) table_of[i] = -1;
    int slot = 0;
    for (int a = 0; a < ns; ++a) {
        const int selector = (int)cur.u8();
        const unsigned td_ta = cur.u8();
        const int td = (int)(td_ta >> 4), ta = (int)(td_ta & 15);
        log.log("\tc_id: %d, dc: %d, ac: %d\n", selector, td, ta);
        int ci = -1;
        for (int i = 0; i < plan.num_comp; ++i)
            if (plan.comps[i].id == selector) {
                ci = i;
                break;
            }
        if (ci < 0) return JPEGGPU_INVALID_JPEG;
        // T.81 A.2: scan components follow frame order
        if (a > 0 && ci <= g.frame_comp[a - 1]) return JPEGGPU_INVALID_JPEG;
        if (st.comp_coded[ci]) return JPEGGPU_INVALID_JPEG; // baseline codes a component once
        if (td > 3 || ta > 3) return JPEGGPU_INVALID_JPEG;
        const int gdc = td * 2 + 0, gac = ta * 2 + 1;
        if (!st.huff[gdc].defined || !st.huff[gac].defined) return JPEGGPU_INVALID_JPEG;
        const FrameComp& fc = plan.comps[ci];
        if (!plan.q_defined[fc.qidx]) return JPEGGPU_INVALID_JPEG;
        st.comp_coded[ci] = true;
        if (table_of[gdc] < 0) {
            table_of[gdc] = g.num_tables;
            sp.tables[g.num_tables++] = st.huff[gdc];
        }
        if (table_of[gac] < 0) {
            table_of[gac] = g.num_tables;
            sp.tables[g.num_tables++] = st.huff[gac];
        }
        g.frame_comp[a] = ci;
        sp.qidx[a] = fc.qidx;
        g.ssx[a] = interleaved ? fc.ssx : 1;
        g.ssy[a] = interleaved ? fc.ssy : 1;
        g.du_off[a] = slot;
        g.size_x[a] = fc.size_x;
        g.size_y[a] = fc.size_y;
        // T.81 A.2.3/A.2.4: the coded plane is padded to whole MCUs
        g.blocks_x[a] = ceil_div(fc.size_x, 8 * g.ssx[a]) * g.ssx[a];
        g.blocks_y[a] = ceil_div(fc.size_y, 8 * g.ssy[a]) * g.ssy[a];
        g.mcus_x = g.blocks_x[a] / g.ssx[a];
        g.mcus_y = g.blocks_y[a] / g.ssy[a];
        const int dus = g.ssx[a] * g.ssy[a];
        if (slot + dus > kMaxDuPerMcu) return JPEGGPU_INVALID_JPEG; // T.81 B.2.3
        for (int j = 0; j < dus; ++j) {
            g.slot_dc[slot] = (uint8_t)table_of[gdc];
            g.slot_ac[slot] = (uint8_t)table_of[gac];
            g.slot_comp[slot] = (uint8_t)a;
            ++slot;
        }
    }
    g.dus_per_mcu = slot;
    // MCU counts of all components must agree (they do by construction of A.1.1 sizes, but a
    // non-conforming frame header could break it): recompute from the first and verify.
    for (int a = 0; a < ns; ++a) {
        if (g.blocks_x[a] / g.ssx[a] != g.mcus_x || g.blocks_y[a] / g.ssy[a] != g.mcus_y)
            return JPEGGPU_INVALID_JPEG;
    }
    const int64_t total = (int64_t)g.mcus_x * g.mcus_y;
    if (total * g.dus_per_mcu * 64 >= (1ll << 31)) return JPEGGPU_NOT_SUPPORTED;
    g.total_mcus = (int)total;
    g.restart_interval = plan.restart_interval ? plan.restart_interval : g.total_mcus;
    g.du_prefix[0] = 0;
    for (int a = 0; a < ns; ++a)
        g.du_prefix[a + 1] = g.du_prefix[a] + (uint32_t)(g.blocks_x[a] * g.blocks_y[a]);
    cur.u8(); // Ss, Se, Ah/Al: fixed for baseline (src/reader.cpp:427-429 ignores them too)
    cur.u8();
    cur.u8();
    ++plan.num_scans;
    // Deferred walk (parser.hpp): the one scan of the file, no DRI, EOI at the very end.
    const size_t left = cur.left();
    if ((walk == kWalkDevice || walk == kWalkDeviceAlways) && plan.num_scans == 1 && ns == plan.num_comp &&
        left >= (walk == kWalkDevice ? kDeferMinBytes : 2) && left < (1ull << 31) - 4096) {
        // the scan is planned as if it filled the rest of the file; T.81 E.1.4: ceil(MCUs / Ri)
        // restart segments (+ slack for stray markers in damaged files)
        sp.begin = (size_t)(cur.p - cur.begin);
        sp.end = (size_t)(cur.end - cur.begin);
        g.scan_bytes = (uint32_t)(sp.end - sp.begin);
        g.scan_avail = g.scan_bytes;
        const int64_t expected = (total + g.restart_interval - 1) / g.restart_interval;
        if (expected + kDeferSegSlack < (1ll << 28)) {
            g.num_segments = (int)expected + kDeferSegSlack;
            g.num_subs = (g.scan_bytes + (uint32_t)kSubBytes - 1u) / (uint32_t)kSubBytes + (uint32_t)g.num_segments;
            plan.walk_pending = true;
            plan.device_walk = true;
            return JPEGGPU_SUCCESS;
        }
    }
    if (walk == kWalkDeferred && plan.num_scans == 1 && ns == plan.num_comp && plan.restart_interval == 0 &&
        left >= kDeferMinBytes + 2 && left - 2 < (1ull << 31) - 4096 && cur.end[-2] == 0xff &&
        cur.end[-1] == 0xd9) {
        sp.begin = (size_t)(cur.p - cur.begin);
        sp.end = (size_t)(cur.end - 2 - cur.begin);
        g.scan_bytes = (uint32_t)(sp.end - sp.begin);
        g.scan_avail = g.scan_bytes + 2u;
        g.num_segments = 1 + kDeferSegSlack;
        g.num_subs = (g.scan_bytes + (uint32_t)kSubBytes - 1u) / (uint32_t)kSubBytes + (uint32_t)kDeferSegSlack;
        plan.walk_pending = true;
        return JPEGGPU_SUCCESS;
    }
    return walk_scan(cur, sp);
}

// The marker loop from the current position to EOI (or to a deferred scan).
jpeggpu_status marker_loop(Cursor& cur, ImagePlan& plan, const Logger& log, int walk)
{
    ParseState& st = plan.st;
    for (;;) {
        if (cur.left() < 2) return JPEGGPU_INVALID_JPEG;
        if (cur.u8() != 0xff) return JPEGGPU_INVALID_JPEG;
        unsigned m = cur.u8();
        while (m == 0xff && cur.left() > 0) m = cur.u8(); // fill bytes before a marker
        log.log("%s\n", marker_name(m));
        jpeggpu_status s = JPEGGPU_SUCCESS;
        if (m == 0xc0 || m == 0xc1) {
            if (st.found_sof) return JPEGGPU_INVALID_JPEG;
            st.found_sof = true;
            s = read_sof(cur, plan, log);
        } else if (m >= 0xc2 && m <= 0xcf && m != 0xc4 && m != 0xc8 && m != 0xcc) {
            log.log("\tunsupported JPEG type\n");
            return JPEGGPU_NOT_SUPPORTED;
        } else if (m == 0xc4) {
            s = read_dht(cur, st, log);
        } else if (m == 0xd9) {
            break;
        } else if (m == 0xda) {
            s = read_sos(cur, plan, st, log, walk);
            if (s == JPEGGPU_SUCCESS && plan.walk_pending) return s; // resumed by finish_deferred_walk
        } else if (m == 0xdb) {
            s = read_dqt(cur, plan, log);
        } else if (m == 0xdd) {
            s = read_dri(cur, plan, log);
        } else {
            s = skip_segment(cur, log);
        }
        if (s != JPEGGPU_SUCCESS) return s;
    }
    if (!st.found_sof) return JPEGGPU_INVALID_JPEG;
    for (int c = 0; c < plan.num_comp; ++c)
        if (!st.comp_coded[c]) {
            log.log("\tcomponent with index %d not defined in scan\n", c);
            return JPEGGPU_INVALID_JPEG;
        }
    return JPEGGPU_SUCCESS;
}

jpeggpu_status parse_impl(const uint8_t* data, size_t size, ImagePlan& plan, const Logger& log,
                          int walk)
{
    Cursor cur{data, data, data + size};
    plan.st = ParseState{};
    plan.walk_pending = false;
    plan.device_walk = false;
    ++plan.generation;
    plan.width = plan.height = plan.num_comp = 0;
    plan.restart_interval = 0;
    plan.num_scans = 0;
    memset(plan.q_defined, 0, sizeof(plan.q_defined));
    plan.file = data;
    plan.file_size = size;

    if (cur.left() < 2) return JPEGGPU_INVALID_JPEG;
    if (cur.u8() != 0xff || cur.u8() != 0xd8) return JPEGGPU_INVALID_JPEG;
    log.log("marker SOI\n");
    return marker_loop(cur, plan, log, walk);
}

jpeggpu_status finish_impl(ImagePlan& plan, const Logger& log, WalkPool* pool)
{
    ScanPlan& sp = plan.scans[0];
    const ScanGeom bound = sp.geom;
    Cursor cur{plan.file + sp.begin, plan.file, plan.file + plan.file_size};
    plan.walk_pending = false;
    plan.device_walk = false;
    jpeggpu_status s = JPEGGPU_SUCCESS;
    if (!walk_parallel(pool, plan, cur, sp, &s)) {
        cur = Cursor{plan.file + sp.begin, plan.file, plan.file + plan.file_size};
        sp.segs.clear();
        sp.tiles.clear();
        s = walk_scan(cur, sp);
    }
    if (s != JPEGGPU_SUCCESS) return s;
    // the scratch plan was sized from the bounds
    if (sp.geom.scan_bytes > bound.scan_bytes || sp.geom.scan_avail > bound.scan_avail ||
        sp.geom.num_subs > bound.num_subs || sp.geom.num_segments > bound.num_segments)
        return JPEGGPU_INVALID_JPEG; // more stray restart markers than the slack allows
    return marker_loop(cur, plan, log, kWalkEager);
}

} // namespace

jpeggpu_status parse_jpeg(const uint8_t* data, size_t size, ImagePlan& plan, const Logger& log,
                          int walk)
{
    try {
        return parse_impl(data, size, plan, log, walk);
    } catch (const std::bad_alloc&) {
        return JPEGGPU_OUT_OF_HOST_MEMORY;
    }
}

jpeggpu_status finish_deferred_walk(ImagePlan& plan, const Logger& log, WalkPool* pool)
{
    if (!plan.walk_pending) return JPEGGPU_SUCCESS;
    try {
        return finish_impl(plan, log, pool);
    } catch (const std::bad_alloc&) {
        plan.walk_pending = false;
        return JPEGGPU_OUT_OF_HOST_MEMORY;
    }
}

} // namespace jb2
