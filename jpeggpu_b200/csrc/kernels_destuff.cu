// kernels_destuff.cu — byte unstuffing + restart-segment indexing, entirely on the device
// (reference stage a1, src/decode_destuff.cu:37-148 + the three CUB scans at :220-316, AND the
// host walk that feeds it, src/reader.cpp:443-489).
//
// One launch, one pass, 16-byte vector loads.  Each CTA takes a 4 KiB tile of stuffed bytes:
// classify (data / stuffed 0xff00 / RSTn / fill / end of scan), block scan, compaction through
// shared memory, coalesced stores.  What a tile needs to know about the bytes BEFORE it - how
// many subsequences the closed restart segments occupy, how many bytes the open segment already
// holds, how many segments were closed, whether the scan has ended - is a prefix over per-tile
// aggregates {closings, bytes before the first closing, subsequences of the segments closed inside,
// bytes after the last closing, ended}.  The combine is associative (not commutative), so the
// prefix is a single-pass decoupled look-back: a tile publishes its aggregate, reads its
// predecessors back to the nearest published inclusive prefix, publishes its own.  Tiles are
// numbered by an atomic ticket, so a tile only ever waits for tiles that are already running.
// The kernel also writes the segment directory (one SegDesc per restart segment), the zero padding
// behind every segment, and - from the tile that holds the end of the scan - the totals every
// later kernel reads (ScanInfo: subsequences, segments): the host never looks at the entropy-coded
// bytes.  It also clears the control block and the sync mailboxes of the Huffman stage.
// Output is the destuffed scan stored as big-endian 32-bit words (byte d lands at d ^ 3) so the
// Huffman kernels read MSB-first words directly; segment k starts at segs[k].sub_off*kSubBytes.
// Algorithmic bytes: Sb read + Sd written + 16 B per tile of look-back traffic.
#include "kernels.hpp"

namespace jb2 {
namespace {

constexpr uint32_t kMaxClose = kDestuffTile / 2 + 1; // closings (RSTn / end of scan) in one tile

struct Chunk {
    uint32_t keep; // bit i: byte i is a data byte
    uint32_t rst;  // bit i: byte i is the second byte of an RSTn marker
    uint32_t term; // index of the first byte that ends the scan (16: none)
    uint8_t val[kDestuffBytesPerThread]; // emitted value of byte i (0xff for a stuffed 0x00)
};

// The per-byte rule of src/decode_destuff.cu:37-44 and :64-71, plus what the reference's host
// walk decides (src/reader.cpp:450-487): 0xff followed by anything but 0x00, RSTn or another
// 0xff (fill) ends the scan.
// lead: bytes at the front of src that precede the scan (never data; the scan starts cold).
__device__ __forceinline__ void classify(const uint8_t* __restrict__ src, uint32_t n, uint32_t lead,
                                         uint32_t base, Chunk& ck)
{
    ck.keep = 0;
    ck.rst = 0;
    ck.term = 16;
    if (base >= n) return;
    uint32_t w[4];
    if (base + 16 <= n) {
        const uint4 v = *reinterpret_cast<const uint4*>(src + base);
        w[0] = v.x, w[1] = v.y, w[2] = v.z, w[3] = v.w;
    } else {
        w[0] = w[1] = w[2] = w[3] = 0;
        for (uint32_t i = 0; base + i < n; ++i)
            w[i >> 2] |= (uint32_t)src[base + i] << (8 * (i & 3));
    }
    uint32_t prev = base ? src[base - 1] : 0u;
    const uint32_t count = min(16u, n - base);
    uint32_t term = 16;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const uint32_t b = (w[i >> 2] >> (8 * (i & 3))) & 0xffu;
        const bool in = (uint32_t)i < count && base + (uint32_t)i >= lead;
        const bool after_ff = prev == 0xffu && base + (uint32_t)i > lead;
        const bool keep = in && (after_ff ? b == 0u : b != 0xffu);
        const bool rst = in && after_ff && b >= 0xd0u && b <= 0xd7u;
        const bool end = in && after_ff && b != 0u && b != 0xffu && !rst;
        ck.keep |= (uint32_t)keep << i;
        ck.rst |= (uint32_t)rst << i;
        term = (end && term == 16u) ? (uint32_t)i : term;
        ck.val[i] = after_ff ? 0xffu : (uint8_t)b;
        prev = b;
    }
    ck.term = term;
}

// Inclusive block scan of a packed (kept | rst << 16) count (< 2^31); returns the exclusive value.
// any_flag: whether `flag` holds in any thread of the block - it travels in bit 31 of the warp
// sums, so the question costs no barrier of its own.
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, bool flag, uint32_t* warp_sums,
                                                         uint32_t& block_total, bool& any_flag)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += up;
    }
    const uint32_t warp_flag = __any_sync(0xffffffffu, flag) ? 0x80000000u : 0u;
    if (lane == 31) warp_sums[warp] = inc | warp_flag;
    __syncthreads();
    uint32_t before = 0, total = 0, flags = 0;
#pragma unroll
    for (int w = 0; w < kDestuffThreads / 32; ++w) {
        const uint32_t s = warp_sums[w];
        if (w < warp) before += s & 0x7fffffffu;
        total += s & 0x7fffffffu;
        flags |= s;
    }
    block_total = total;
    any_flag = (flags >> 31) != 0u;
    return before + inc - v;
}

// ---- the prefix element and its combine ----------------------------------------------------
struct Elem {
    uint32_t nclose; // segment closings (RSTn markers, the end of the scan)
    uint32_t head;   // data bytes before the first closing (all of them if there is none)
    uint32_t inner;  // subsequences of the segments that begin AND end inside
    uint32_t tail;   // data bytes after the last closing (0 if there is none)
    uint32_t ended;  // the scan has ended
};

__device__ __forceinline__ uint32_t subs_of(uint32_t bytes)
{
    return (bytes + (uint32_t)kSubBytes - 1u) / (uint32_t)kSubBytes;
}

// L covers the bytes before R.  Nothing follows the end of the scan.
__device__ __forceinline__ Elem combine(const Elem& L, const Elem& R)
{
    if (L.ended) return L;
    Elem o;
    o.ended = R.ended;
    if (R.nclose == 0) {
        o.nclose = L.nclose;
        o.inner = L.inner;
        o.head = L.nclose ? L.head : L.head + R.head;
        o.tail = L.nclose ? L.tail + R.head : 0u;
    } else if (L.nclose == 0) {
        o.nclose = R.nclose;
        o.head = L.head + R.head;
        o.inner = R.inner;
        o.tail = R.tail;
    } else {
        o.nclose = L.nclose + R.nclose;
        o.head = L.head;
        o.inner = L.inner + subs_of(L.tail + R.head) + R.inner;
        o.tail = R.tail;
    }
    return o;
}

// Two 64-bit words per tile: {flag:2 | ended:1 | nclose:29 | head:32}, {flag:2 | inner:30 | tail:32}.
// flag 1: the tile's own aggregate, 2: inclusive prefix.  Both words carry the flag, so a reader
// can tell a half-written pair (the flags differ) and reads again.
constexpr unsigned long long kFlagAgg = 1ull << 62, kFlagPrefix = 2ull << 62;

__device__ __forceinline__ void publish(unsigned long long* d, const Elem& e, unsigned long long flag)
{
    const unsigned long long w0 =
        flag | ((unsigned long long)(e.ended & 1u) << 61) | ((unsigned long long)(e.nclose & 0x1fffffffu) << 32) | e.head;
    const unsigned long long w1 = flag | ((unsigned long long)(e.inner & 0x3fffffffu) << 32) | e.tail;
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(d), "l"(w0) : "memory");
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(d + 1), "l"(w1) : "memory");
}

// (one 16-byte load; it may tear, which the flags in both words reveal)
__device__ __forceinline__ void load_desc(const unsigned long long* d, unsigned long long& w0,
                                          unsigned long long& w1)
{
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(d) : "memory");
}
// Returns the flag (0: not ready or half written).
__device__ __forceinline__ uint32_t unpack_desc(unsigned long long w0, unsigned long long w1, Elem& e)
{
    const uint32_t f0 = (uint32_t)(w0 >> 62), f1 = (uint32_t)(w1 >> 62);
    e.ended = (uint32_t)(w0 >> 61) & 1u;
    e.nclose = (uint32_t)(w0 >> 32) & 0x1fffffffu;
    e.head = (uint32_t)w0;
    e.inner = (uint32_t)(w1 >> 32) & 0x3fffffffu;
    e.tail = (uint32_t)w1;
    return f0 == f1 ? f0 : 0u;
}

__device__ __forceinline__ Elem shfl_down_elem(const Elem& e, int d)
{
    Elem o;
    const uint32_t ce = __shfl_down_sync(0xffffffffu, e.nclose | (e.ended << 31), d);
    o.nclose = ce & 0x7fffffffu;
    o.ended = ce >> 31;
    o.head = __shfl_down_sync(0xffffffffu, e.head, d);
    o.inner = __shfl_down_sync(0xffffffffu, e.inner, d);
    o.tail = __shfl_down_sync(0xffffffffu, e.tail, d);
    return o;
}

// The combine of all tiles before `tile`: decoupled look-back by the whole CTA.  All tiles of an
// image are resident at once and start together, so when a tile looks back its predecessors have
// published aggregates but hardly any inclusive prefix: the look-back usually runs all the way to
// the start of the scan.  Every thread therefore takes kLookK descriptors of a window of
// kDestuffThreads * kLookK tiles and has all its loads in flight at once: one round trip to L2 and
// one block-wide reduction cover 1024 tiles (every tile of a 4 MiB scan).  Thread t, element j
// looks at distance t*kLookK + j behind `tile`: a thread's elements are consecutive tiles, so it
// combines them itself before the warps reduce.
// Called by every thread; the result is valid in every thread.
constexpr int kLookK = 4;
struct LookBackSmem {
    Elem warp_elem[kDestuffThreads / 32];
    uint32_t warp_stop[kDestuffThreads / 32];
};

__device__ __forceinline__ Elem look_back(const unsigned long long* desc, uint32_t tile, LookBackSmem& sm)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kWarps = kDestuffThreads / 32;
    constexpr uint32_t kWindow = (uint32_t)kDestuffThreads * kLookK;
    const Elem none = {0u, 0u, 0u, 0u, 0u};
    Elem carry = none; // combine of the tiles between the window and `tile`
    for (uint32_t base = 0; base < tile; base += kWindow) {
        Elem e[kLookK];
        uint32_t dist[kLookK];
        uint32_t pending = 0, nearest = 0xffffffffu; // nearest distance holding an inclusive prefix
#pragma unroll
        for (int j = 0; j < kLookK; ++j) {
            dist[j] = base + threadIdx.x * (uint32_t)kLookK + (uint32_t)j;
            e[j] = none;
            if (dist[j] < tile) pending |= 1u << j;
            else if (dist[j] == tile) nearest = dist[j]; // the (empty) prefix before the scan
        }
        // every earlier tile is running or done (tickets); the bound only matters if the
        // descriptors were not zero when the kernel started (decode without transfer): the
        // decode is then garbage, but it ends
        for (int spin = 0; pending && spin < (1 << 16); ++spin) {
            // all loads first: a test of a loaded flag between them would serialise the round trips
            unsigned long long w0[kLookK], w1[kLookK];
#pragma unroll
            for (int j = 0; j < kLookK; ++j) {
                w0[j] = w1[j] = 0ull;
                if ((pending >> j) & 1u) load_desc(desc + 2ull * (tile - 1u - dist[j]), w0[j], w1[j]);
            }
#pragma unroll
            for (int j = 0; j < kLookK; ++j) {
                if (!((pending >> j) & 1u)) continue;
                Elem got;
                const uint32_t flag = unpack_desc(w0[j], w1[j], got);
                if (flag == 0u) continue;
                e[j] = got;
                pending &= ~(1u << j);
                if (flag == 2u) nearest = min(nearest, dist[j]);
            }
            if (pending) __nanosleep(32);
        }
        if (pending) { // gave up: end the scan here
#pragma unroll
            for (int j = 0; j < kLookK; ++j)
                if ((pending >> j) & 1u) {
                    e[j] = Elem{0u, 0u, 0u, 0u, 1u};
                    nearest = min(nearest, dist[j]);
                }
        }
        // Nothing beyond the nearest inclusive prefix matters.  Every warp reduces its own 128
        // tiles first - cut at ITS nearest prefix, which can only be farther than the block's -
        // so that ONE barrier suffices: afterwards the warps that lie entirely beyond the block's
        // nearest prefix are simply left out.
        nearest = __reduce_min_sync(0xffffffffu, nearest);
        uint32_t busy = 0, heads = 0;
#pragma unroll
        for (int j = 0; j < kLookK; ++j) {
            if (dist[j] > nearest) e[j] = none;
            busy |= e[j].nclose | e[j].ended;
            heads += e[j].head;
        }
        Elem mine = Elem{0u, __reduce_add_sync(0xffffffffu, heads), 0u, 0u, 0u};
        if (__any_sync(0xffffffffu, busy != 0u)) {
            // a closing among the warp's tiles (restart markers, the end of the scan): ordered
            // reduction (a higher element, lane or warp = a larger distance = an earlier tile): the
            // thread's own elements, then a tree over the lanes: lane 0 ends with E[31] o ... o E[0]
            mine = e[0];
#pragma unroll
            for (int j = 1; j < kLookK; ++j) mine = combine(e[j], mine);
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const Elem far = shfl_down_elem(mine, d);
                if (lane + d < 32) mine = combine(far, mine);
            }
        }
        if (lane == 0) {
            sm.warp_elem[warp] = mine;
            sm.warp_stop[warp] = nearest;
        }
        __syncthreads();
        uint32_t stop = 0xffffffffu;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) stop = min(stop, sm.warp_stop[w]);
        // warps whose first tile lies beyond the stop do not count (the one that holds it cut itself there)
        Elem win = none;
#pragma unroll
        for (int w = kWarps - 1; w >= 0; --w) {
            const uint32_t warp_first = base + (uint32_t)(w * 32 * kLookK);
            if (warp_first <= stop) win = combine(win, sm.warp_elem[w]);
        }
        carry = combine(win, carry);
        if (stop != 0xffffffffu) break; // an inclusive prefix was found (the start of the scan counts as one)
        __syncthreads();                // the next window reuses the shared arrays
    }
    return carry;
}

struct DestuffArgs {
    const uint8_t* src;        // 16-byte aligned; the scan starts `lead` bytes in
    uint32_t lead;
    uint32_t n;                // bytes at src (the scan may end earlier, at its terminating marker)
    uint8_t* dst;              // destuffed scan
    SegDesc* segs;             // out: segment directory
    uint32_t seg_cap, sub_cap; // capacities of the directory / of dst in subsequences
    unsigned long long* desc;  // look-back descriptors, 2 words per tile; zero on entry
    uint32_t* ticket;          // zero on entry
    ScanInfo* info;            // out
    uint32_t* clear;           // control block + mailboxes of the later stages
    uint32_t clear_words;
};

__device__ __forceinline__ void destuff_body(const DestuffArgs& a, const uint32_t block, const uint32_t nblocks)
{
    __shared__ uint32_t warp_sums[kDestuffThreads / 32];
    __shared__ uint8_t stage[kDestuffTile];
    __shared__ uint16_t seg_of[kDestuffTile];  // closings before the staged byte, inside the tile
    __shared__ uint32_t close_at[kMaxClose];   // data bytes of the tile before closing i
    __shared__ uint32_t sub_rel[kMaxClose + 1]; // subsequences of the segments closed by closings 1..i-1
    __shared__ uint32_t s_tile, s_term, s_flags, s_inner;
    __shared__ LookBackSmem s_look;

    // first kernel of a scan (launched behind the upload copies): let huff_sync start its
    // prologue (tables, zero-fill of the coefficient buffer) right away
    grid_dependency_wait();
    grid_dependency_trigger();
#ifdef JB2_TIMELINE
    Control* const tl_ctrl = a.clear ? reinterpret_cast<Control*>(a.clear - kTimelineWords) : nullptr;
#endif
    JB2_TL(tl_ctrl, 0, true);
    for (uint32_t i = block * kDestuffThreads + threadIdx.x; i < a.clear_words;
         i += nblocks * kDestuffThreads)
        a.clear[i] = 0u;
    if (threadIdx.x == 0) {
        s_tile = atomicAdd(a.ticket, 1u);
        s_term = kDestuffTile;
        s_flags = 0u;
    }
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile >= nblocks) return; // the ticket was not zero on entry (decode without transfer)
    JB2_TL(tl_ctrl, 16, false); // latest ticket
    const uint32_t n = a.n;
    const uint32_t tile_begin = tile * (uint32_t)kDestuffTile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    const uint32_t base = tile_begin + threadIdx.x * kDestuffBytesPerThread;
    Chunk ck;
    classify(a.src, n, a.lead, base, ck);
    // The scan ends at its first terminating marker or, failing that, with the data: nothing at or
    // behind that position counts.
    const bool last_tile = tile_begin + (uint32_t)kDestuffTile >= n;
    // Nearly every tile has no terminating marker: scan its counts right away and learn from the
    // same barrier whether that was true; only a tile that holds one (or the last tile) trims its
    // masks and scans again.
    uint32_t tile_total = 0, ex = 0;
    bool ends = last_tile; // block-uniform
    if (!ends) {
        ex = block_exclusive_scan(__popc(ck.keep) | (__popc(ck.rst) << 16), ck.term < 16u, warp_sums, tile_total,
                                  ends);
        if (ends) __syncthreads(); // warp_sums is written again below
    }
    if (ends) {
        if (ck.term < 16u) atomicMin(&s_term, threadIdx.x * kDestuffBytesPerThread + ck.term);
        __syncthreads();
        uint32_t term = s_term;
        if (term == (uint32_t)kDestuffTile && last_tile) {
            term = n - tile_begin; // 1 .. kDestuffTile: the data ran out before a marker was seen
            if (threadIdx.x == 0) s_flags = kScanUnterminated;
        }
        const uint32_t mine = threadIdx.x * kDestuffBytesPerThread;
        const uint32_t live = term <= mine ? 0u : term - mine >= 16u ? 0xffffu : (1u << (term - mine)) - 1u;
        ck.keep &= live;
        ck.rst &= live;
        if (threadIdx.x == 0) s_term = term;
        bool unused;
        ex = block_exclusive_scan(__popc(ck.keep) | (__popc(ck.rst) << 16), false, warp_sums, tile_total, unused);
    }
    uint32_t pos = ex & 0xffffu, seg = ex >> 16;
    const uint32_t tile_kept = tile_total & 0xffffu;
    const bool ends_here = s_term < (uint32_t)kDestuffTile || (last_tile && s_term == n - tile_begin);
    uint32_t nclose = tile_total >> 16;
    unsigned long long* mine = a.desc + 2ull * tile;
    Elem agg;
    if (nclose == 0 && !ends_here) { // block-uniform
        // Nothing closes inside the tile (every tile but the last of a scan without DRI, most tiles
        // otherwise): the aggregate is the byte count, known right after the block scan.  Publish
        // it BEFORE the bytes are staged - the tiles behind this one are waiting for it.
        agg = Elem{0u, tile_kept, 0u, 0u, 0u};
        if (threadIdx.x == 0 && tile != 0) publish(mine, agg, kFlagAgg);
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if ((ck.keep >> i) & 1u) stage[pos++] = ck.val[i];
        if (tile == 0) __syncthreads(); // (the look-back of every other tile has a barrier)
    } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            if ((ck.keep >> i) & 1u) {
                stage[pos] = ck.val[i];
                seg_of[pos] = (uint16_t)seg;
                ++pos;
            }
            if ((ck.rst >> i) & 1u) close_at[seg++] = pos;
        }
        if (ends_here) { // the end of the scan closes the open segment
            if (threadIdx.x == 0) close_at[nclose] = tile_kept;
            ++nclose;
        }
        __syncthreads();

        // subsequences of the segments closed inside the tile by closings 1 .. nclose-1, as a running
        // sum (restart markers are rare - one per MCU row in camera files - so one warp does it)
        if (nclose <= 1) {
            if (threadIdx.x < 2) sub_rel[threadIdx.x] = 0u;
            if (threadIdx.x == 0) s_inner = 0u;
        } else if (warp == 0) {
            uint32_t run = 0;
            for (uint32_t i0 = 1;; i0 += 32u) {
                const uint32_t i = i0 + (uint32_t)lane;
                uint32_t v = i < nclose ? subs_of(close_at[i] - close_at[i - 1]) : 0u, inc = v;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t up = __shfl_up_sync(0xffffffffu, inc, d);
                    if (lane >= d) inc += up;
                }
                if (i <= nclose) sub_rel[i] = run + inc - v;
                run += __shfl_sync(0xffffffffu, inc, 31);
                if (i0 + 31u >= nclose) break; // the chunk held index nclose: `run` is the aggregate's `inner`
            }
            if (lane == 0) s_inner = run;
        }
        __syncthreads();
        agg.nclose = nclose;
        agg.head = close_at[0];
        agg.inner = s_inner;
        agg.tail = tile_kept - close_at[nclose - 1];
        agg.ended = ends_here ? 1u : 0u;
        if (threadIdx.x == 0 && tile != 0) publish(mine, agg, kFlagAgg);
    }
    JB2_TL(tl_ctrl, 17, false); // latest aggregate
    JB2_TL(tl_ctrl, 18, true);  // earliest aggregate
    const Elem before = tile ? look_back(a.desc, tile, s_look) : Elem{0u, 0u, 0u, 0u, 0u};
    JB2_TL(tl_ctrl, 19, false); // latest look-back done
    if (threadIdx.x == 0) {
        const Elem incl = combine(before, agg);
        publish(mine, incl, kFlagPrefix);
        if (ends_here && !before.ended) {
            // the end of the scan: totals for the later stages (the end closed the last segment)
            const uint32_t subs = subs_of(incl.head) + incl.inner;
            ScanInfo si;
            si.num_subs = min(subs, a.sub_cap);
            si.num_segments = min(incl.nclose, a.seg_cap);
            si.scan_end = tile_begin + s_term - a.lead;
            si.status = s_flags | ((subs > a.sub_cap || incl.nclose > a.seg_cap) ? kScanOverflow : 0u);
            *a.info = si;
        }
    }
    if (before.ended) return; // bytes behind the end of the scan
    const uint32_t closed0 = before.nclose ? subs_of(before.head) + before.inner : 0u; // subsequences
    const uint32_t open0 = before.nclose ? before.tail : before.head; // bytes already in the open segment
    const uint32_t segs0 = before.nclose;
    const unsigned long long cap_bytes = (unsigned long long)a.sub_cap * kSubBytes;

    // the segments this tile closes: directory entry + zero padding to a whole subsequence
    const uint32_t first_subs = nclose ? subs_of(open0 + close_at[0]) : 0u;
    for (uint32_t i = threadIdx.x; i < nclose; i += kDestuffThreads) {
        const uint32_t nbytes = i ? close_at[i] - close_at[i - 1] : open0 + close_at[0];
        const uint32_t sub_off = i ? closed0 + first_subs + sub_rel[i] : closed0;
        const uint32_t g = segs0 + i;
        if (g >= a.seg_cap || (unsigned long long)sub_off + subs_of(nbytes) > a.sub_cap) continue;
        SegDesc sd;
        sd.sub_off = sub_off;
        sd.nbytes = nbytes;
        sd.kept_before = 0;
        sd.src_begin = 0;
        a.segs[g] = sd;
        const uint32_t d_end = sub_off * (uint32_t)kSubBytes + nbytes;
        const uint32_t pad_end = (d_end + (uint32_t)kSubBytes - 1u) / (uint32_t)kSubBytes * (uint32_t)kSubBytes;
        for (uint32_t z = d_end; z < pad_end; ++z) a.dst[z ^ 3u] = 0;
    }

    if (nclose == 0 || close_at[0] == tile_kept) {
        // Every data byte of the tile belongs to the segment that was open on entry (every tile of
        // a scan without DRI, most tiles otherwise): one contiguous run of output bytes.  Whole
        // big-endian words are assembled from the staged bytes and stored with 4-byte stores;
        // only the run's unaligned head and tail go byte by byte.
        const unsigned long long d0l = (unsigned long long)closed0 * kSubBytes + open0;
        if (d0l + tile_kept > cap_bytes) return; // a damaged stream outgrew the plan (flagged above)
        const uint32_t d0 = (uint32_t)d0l;
        const uint32_t d_end = d0 + tile_kept;
        const uint32_t w_begin = (d0 + 3u) & ~3u, w_end = d_end & ~3u; // whole words [w_begin, w_end)
        if (w_begin < w_end) {
            uint32_t* words = reinterpret_cast<uint32_t*>(a.dst);
            for (uint32_t w = w_begin + 4u * threadIdx.x; w < w_end; w += 4u * kDestuffThreads) {
                const uint32_t o = w - d0;
                words[w >> 2] = ((uint32_t)stage[o] << 24) | ((uint32_t)stage[o + 1] << 16) |
                                ((uint32_t)stage[o + 2] << 8) | (uint32_t)stage[o + 3];
            }
        }
        // head [d0, min(w_begin, d_end)) and tail [max(w_end, head end), d_end): < 8 bytes
        const uint32_t head_end = w_begin < d_end ? w_begin : d_end;
        const uint32_t tail_begin = w_end > head_end ? w_end : head_end;
        const uint32_t edge = (head_end - d0) + (d_end - tail_begin);
        if (threadIdx.x < edge) {
            const uint32_t d = threadIdx.x < head_end - d0 ? d0 + threadIdx.x
                                                           : tail_begin + (threadIdx.x - (head_end - d0));
            a.dst[d ^ 3u] = stage[d - d0];
        }
        return;
    }
    for (uint32_t o = threadIdx.x; o < tile_kept; o += kDestuffThreads) {
        const uint32_t s = seg_of[o];
        unsigned long long d;
        if (s == 0) d = (unsigned long long)closed0 * kSubBytes + open0 + o;
        else d = (unsigned long long)(closed0 + first_subs + sub_rel[s]) * kSubBytes + (o - close_at[s - 1]);
        if (d < cap_bytes && segs0 + s < a.seg_cap) a.dst[(uint32_t)d ^ 3u] = stage[o];
    }
}

DestuffArgs destuff_args(const ScanGeom& g, const ScanBufs& b)
{
    DestuffArgs a;
    a.src = b.scan_bytes;
    a.lead = b.scan_lead;
    a.n = b.scan_lead + g.scan_avail;
    a.dst = reinterpret_cast<uint8_t*>(b.destuffed);
    a.segs = b.segs;
    a.seg_cap = (uint32_t)g.num_segments;
    a.sub_cap = g.num_subs;
    a.desc = b.destuff_desc;
    a.ticket = b.destuff_ticket;
    a.info = b.info;
    // the control block, the mailboxes and the DC flags are contiguous (planner): one clear range
    a.clear = reinterpret_cast<uint32_t*>(b.ctrl) + kTimelineWords;
    a.clear_words = b.clear_words - kTimelineWords;
    return a;
}

// (six CTAs per SM: every tile of a 12 MP scan is resident at once - a tile of a second wave would
// spend the whole first wave waiting for its ticket)
__global__ void __launch_bounds__(kDestuffThreads, 6)
destuff_kernel(const DestuffArgs a)
{
    destuff_body(a, blockIdx.x, gridDim.x);
    JB2_TL(reinterpret_cast<Control*>(a.clear - kTimelineWords), 1, false);
}

__global__ void __launch_bounds__(kDestuffThreads, 6)
destuff_batch_kernel(const BatchImage* __restrict__ images)
{
    const BatchImage& im = images[blockIdx.y];
    const uint32_t nblocks = im.ctas[kBkDestuff];
    if (blockIdx.x >= nblocks) return;
    __shared__ DestuffArgs a;
    if (threadIdx.x == 0) {
        DestuffArgs t;
        t.src = im.b.scan_bytes;
        t.lead = im.b.scan_lead;
        t.n = im.b.scan_lead + im.g.scan_avail;
        t.dst = reinterpret_cast<uint8_t*>(im.b.destuffed);
        t.segs = im.b.segs;
        t.seg_cap = (uint32_t)im.g.num_segments;
        t.sub_cap = im.g.num_subs;
        t.desc = im.b.destuff_desc;
        t.ticket = im.b.destuff_ticket;
        t.info = im.b.info;
        t.clear = reinterpret_cast<uint32_t*>(im.b.ctrl) + kTimelineWords;
        t.clear_words = im.b.clear_words - kTimelineWords;
        a = t;
    }
    __syncthreads();
    destuff_body(a, blockIdx.x, nblocks);
}

} // namespace

uint32_t destuff_tiles(const ScanGeom& g, const ScanBufs& b)
{
    return (b.scan_lead + g.scan_avail + kDestuffTile - 1) / kDestuffTile;
}

cudaError_t launch_destuff_batch(const BatchLaunch& bl, cudaStream_t stream)
{
    static PerDeviceOnce configured;
    configured.run([] { prefer_max_shared(destuff_batch_kernel); });
    destuff_batch_kernel<<<dim3(bl.max_ctas[kBkDestuff], bl.count), kDestuffThreads, 0, stream>>>(bl.d_images);
    return cudaGetLastError();
}

cudaError_t launch_destuff(const ScanGeom& g, const ScanBufs& b, cudaStream_t stream, StageTimer* tm)
{
    if (g.scan_avail == 0) return cudaSuccess;
    static PerDeviceOnce configured;
    configured.run([] { prefer_max_shared(destuff_kernel); });
    const uint32_t tiles = destuff_tiles(g, b);
    const DestuffArgs a = destuff_args(g, b);
    tm->mark(stream, kIdDestuff);
    destuff_kernel<<<tiles, kDestuffThreads, 0, stream>>>(a);
    return cudaGetLastError();
}

} // namespace jb2

// ---- destuff without any allocation ------------------------------------------------------------
// jpeggpu_b200_destuff_static (include/jpeggpu_b200.h): the destuff kernel for callers that cannot
// call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  No
// counterpart in the reference.  The look-back
// descriptors and the tile ticket - the kernel's scratch - live in a file-scope device array of a
// fixed size, so every GPU that loads the library holds kStaticScratchBytes until the process
// exits.  The array's name is a host-side stand-in in host code: its device address comes from
// cudaGetSymbolAddress, once per device with that device current.  Every call shares the array
// (calls that overlap on the GPU race) and it keeps the last call's contents: the part a call uses
// is cleared on the caller's stream in front of the kernel.
#include <jpeggpu_b200.h>

#include <cstdio>

namespace jb2 {
namespace {

constexpr size_t kStaticMaxScanBytes = JPEGGPU_B200_DESTUFF_STATIC_MAX_BYTES;
constexpr size_t kStaticMaxTiles = (kStaticMaxScanBytes + kDestuffTile - 1) / kDestuffTile;
constexpr size_t kStaticHeaderWords = 32; // 256 bytes: word 0 holds the ticket
constexpr size_t kStaticWords = kStaticHeaderWords + 2 * kStaticMaxTiles;
static_assert(kStaticWords * sizeof(unsigned long long) == 262400, "kStaticScratchBytes: the size the header states");
constexpr int kStaticMaxDevices = 64;

static __device__ __align__(256) unsigned long long g_destuff_static_scratch[kStaticWords];

static_assert(sizeof(jpeggpu_b200_segment) == sizeof(SegDesc), "the directory is handed out as is");
static_assert(sizeof(jpeggpu_b200_scan_info) == sizeof(ScanInfo), "the totals are handed out as they are");

__global__ void __launch_bounds__(kDestuffThreads, 6)
destuff_static_kernel(const DestuffArgs a)
{
    destuff_body(a, blockIdx.x, gridDim.x);
}

std::mutex g_static_lock;
unsigned long long* g_static_ptr[kStaticMaxDevices] = {};

// Device address of the scratch array on the current device (looked up once per device).
cudaError_t static_scratch(unsigned long long** out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kStaticMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> guard(g_static_lock);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_destuff_static_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        prefer_max_shared(destuff_static_kernel);
        g_static_ptr[dev] = static_cast<unsigned long long*>(p);
    }
    *out = g_static_ptr[dev];
    return cudaSuccess;
}

} // namespace
} // namespace jb2

extern "C" enum jpeggpu_status jpeggpu_b200_destuff_static(
    const uint8_t* d_scan, size_t scan_bytes, uint32_t* d_destuffed, size_t max_subsequences,
    struct jpeggpu_b200_segment* d_segments, size_t max_segments, struct jpeggpu_b200_scan_info* d_info,
    cudaStream_t stream)
{
    using namespace jb2;
    if (!d_scan || !d_destuffed || !d_segments || !d_info || scan_bytes == 0 || max_subsequences == 0 ||
        max_segments == 0)
        return JPEGGPU_INVALID_ARGUMENT;
    if ((reinterpret_cast<uintptr_t>(d_scan) | reinterpret_cast<uintptr_t>(d_destuffed) |
         reinterpret_cast<uintptr_t>(d_segments) | reinterpret_cast<uintptr_t>(d_info)) & 15u)
        return JPEGGPU_INVALID_ARGUMENT;
    if (scan_bytes > kStaticMaxScanBytes) {
        std::fprintf(stderr,
                     "jpeggpu_b200_destuff_static: %zu scan bytes exceed the %zu the static scratch is sized for "
                     "(JPEGGPU_B200_DESTUFF_STATIC_MAX_BYTES)\n",
                     scan_bytes, kStaticMaxScanBytes);
        return JPEGGPU_INVALID_ARGUMENT;
    }
    // every destuffed byte offset must fit 32 bits
    if (max_subsequences > 0xffffffffull / kSubBytes || max_segments > 0x1fffffffull) return JPEGGPU_INVALID_ARGUMENT;

    unsigned long long* scratch = nullptr;
    if (static_scratch(&scratch) != cudaSuccess) return JPEGGPU_INTERNAL_ERROR;
    const uint32_t tiles = (uint32_t)((scan_bytes + kDestuffTile - 1) / kDestuffTile);
    // ticket + the descriptors of this call's tiles start from zero
    if (cudaMemsetAsync(scratch, 0, (kStaticHeaderWords + 2 * (size_t)tiles) * sizeof(unsigned long long), stream) !=
        cudaSuccess)
        return JPEGGPU_INTERNAL_ERROR;
    DestuffArgs a;
    a.src = d_scan;
    a.lead = 0;
    a.n = (uint32_t)scan_bytes;
    a.dst = reinterpret_cast<uint8_t*>(d_destuffed);
    a.segs = reinterpret_cast<SegDesc*>(d_segments);
    a.seg_cap = (uint32_t)max_segments;
    a.sub_cap = (uint32_t)max_subsequences;
    a.desc = scratch + kStaticHeaderWords;
    a.ticket = reinterpret_cast<uint32_t*>(scratch);
    a.info = reinterpret_cast<ScanInfo*>(d_info);
    a.clear = nullptr;
    a.clear_words = 0;
    destuff_static_kernel<<<tiles, kDestuffThreads, 0, stream>>>(a);
    return cudaGetLastError() == cudaSuccess ? JPEGGPU_SUCCESS : JPEGGPU_INTERNAL_ERROR;
}
