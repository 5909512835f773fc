// parser.hpp — host-side marker parser and scan planner (CPU only).
// Replaces src/reader.{hpp,cpp} of the reference: same accepted inputs and status codes
// (SURVEY §8b "Errors"), different products: a per-scan geometry block, compact Huffman
// specs (expanded to LUTs by huff_lut.hpp) and a restart-segment directory with byte counts.
#ifndef JB2_PARSER_HPP_
#define JB2_PARSER_HPP_

#include "common.hpp"

#include <jpeggpu/jpeggpu.h>

#include <vector>

namespace jb2 {

struct HuffSpec {
    bool defined = false;
    bool is_dc = false;
    uint8_t counts[16] = {};
    uint8_t vals[256] = {};
};

struct FrameComp {
    int id, ssx, ssy, qidx;
    int size_x, size_y; // output plane size (T.81 A.1.1)
};

struct ScanPlan {
    ScanGeom geom;
    size_t begin = 0, end = 0;   // byte range of the entropy-coded data in the file
    std::vector<SegDesc> segs;   // restart-segment directory
    std::vector<TileDesc> tiles; // destuff tile directory
    HuffSpec tables[kMaxTables]; // the tables this scan uses, by table slot
    int qidx[kMaxComp] = {};     // quantisation table of each scan component
};

// Marker-parser state that outlives one marker: kept in the plan so that a deferred entropy walk
// (below) can resume the marker loop where parse_header left it.
struct ParseState {
    bool found_sof = false;
    HuffSpec huff[kMaxTables]; // index = th * 2 + class, persists across scans (T.81 B.2.4.2)
    bool comp_coded[kMaxComp] = {};
};

// Deferred walk: capacities planned for a scan whose entropy-coded bytes have not been walked yet
// (the walk may find fewer bytes / subsequences, never more without failing the image).
constexpr size_t kDeferMinBytes = 64 * 1024; // below this the walk is too cheap to be worth it
constexpr int kDeferSegSlack = 16;           // stray RSTn markers tolerated in a scan without DRI

struct ImagePlan {
    int width = 0, height = 0, num_comp = 0;
    int ss_max_x = 0, ss_max_y = 0;
    int restart_interval = 0;
    FrameComp comps[kMaxComp] = {};
    uint8_t qtables[4][64] = {}; // natural order
    bool q_defined[4] = {};
    int num_scans = 0;
    ScanPlan scans[kMaxScans];
    const uint8_t* file = nullptr; // borrowed (include/jpeggpu/jpeggpu.h:83 contract)
    size_t file_size = 0;
    ParseState st;
    // true: scans[0] holds an upper-bound geometry (scan_bytes, num_subs, num_segments) and empty
    // directories; finish_deferred_walk() must run before host code uses exact counts.
    bool walk_pending = false;
    // true (with walk_pending): the device finds the end of the scan and the restart segments
    // itself (kernels_destuff.cu); no host walk is needed to decode.
    bool device_walk = false;
    uint64_t generation = 0; // bumped by every parse: identifies the image a helper job belongs to
};

struct Logger {
    bool on = false;
    void log(const char* fmt, ...) const;
};

// Parses the whole file. Never throws; allocation failure -> JPEGGPU_OUT_OF_HOST_MEMORY.
// allow_defer: when the file is one interleaved scan of all frame components without DRI and
// ends with EOI, stop after the scan header with plan.walk_pending set: the scan is taken to run
// to the EOI at the end of the file, its geometry holds upper bounds, and the byte walk (the only
// part of the parse whose cost grows with the file) is left to finish_deferred_walk(), which the
// decoder runs inside jpeggpu_decoder_transfer while the upload of those bytes is in flight.
// walk: kWalkEager, kWalkDeferred (rules above) or kWalkDevice: every file whose first scan holds
// all frame components stops at the scan header, whatever follows; the scan is taken to reach to
// the end of the file and the device locates its real end (the bytes behind it are uploaded too).
// kWalkDeviceAlways: as kWalkDevice without the 64 KiB floor (batched decode: no per-file host walk).
enum WalkMode { kWalkEager = 0, kWalkDeferred = 1, kWalkDevice = 2, kWalkDeviceAlways = 3 };
jpeggpu_status parse_jpeg(const uint8_t* data, size_t size, ImagePlan& plan, const Logger& log,
                          int walk = kWalkEager);

// Helper threads for the deferred walk of large scans.  The walk is memory bound on one core
// (~35 GB/s: the previous upload has evicted the file from the caches), so scans of at least
// kWalkParallelMin bytes are cut into 64 KB blocks: everything but the tail is a pure counting
// job (0xff bytes per destuff tile, valid only while every 0xff is followed by 0x00) that the
// helpers and the calling thread pull from a shared counter; the calling thread walks the tail,
// where the scan ends, with the general code.  Anything unexpected in a counting block (restart
// or fill markers) discards the parallel result and re-walks sequentially.  The pool is owned by
// one decoder; threads start on first use and only ever read file bytes between
// walk_pool_kick() and the return of finish_deferred_walk(), i.e. inside jpeggpu_decoder_transfer.
struct WalkPool;
constexpr size_t kWalkParallelMin = 1u << 20;
WalkPool* walk_pool_create();
void walk_pool_destroy(WalkPool* pool);
uint64_t walk_pool_stitched(const WalkPool* pool); // walks that completed on the parallel path
// Hands the counting blocks of plan.scans[0] to the helpers (no-op for small scans).
void walk_pool_kick(WalkPool* pool, const ImagePlan& plan);
// Withdraws a kicked job that will not be collected (transfer failed after the kick): returns once
// no helper is reading the file any more.
void walk_pool_cancel(WalkPool* pool);

// Walks the deferred scan, replaces the bounds in scans[0] by the exact values and resumes the
// marker loop behind the scan (same statuses as the eager parse).  No-op unless walk_pending.
// pool: optional, after walk_pool_kick() for the same plan.
jpeggpu_status finish_deferred_walk(ImagePlan& plan, const Logger& log, WalkPool* pool = nullptr);

} // namespace jb2
#endif
