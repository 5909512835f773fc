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
};

struct Logger {
    bool on = false;
    void log(const char* fmt, ...) const;
};

// Parses the whole file. Never throws; allocation failure -> JPEGGPU_OUT_OF_HOST_MEMORY.
jpeggpu_status parse_jpeg(const uint8_t* data, size_t size, ImagePlan& plan, const Logger& log);

} // namespace jb2
#endif
