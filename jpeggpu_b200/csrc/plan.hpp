// plan.hpp — scratch layout of one scan, shared by the per-image decoder (decoder.cpp) and the
// batched path (batch.cpp).  Replaces the reference's stack_allocator dry run
// (src/util.hpp:54-84, src/decoder.cpp:120-155): sizes are pure functions of the parsed geometry.
#ifndef JB2_PLAN_HPP_
#define JB2_PLAN_HPP_

#include "huff_lut.hpp"
#include "kernels.hpp"
#include "parser.hpp"

#include <cstring>
#include <vector>

namespace jb2 {

inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

// Coefficient buffers above this size are cleared by huff_write itself (the 126 MB L2 of a B200
// no longer holds the zero-filled lines until the scattered stores arrive).  Measured on B200,
// whole decode in us, up-front fill / in-write fill: 24 MP (73 MB) 174 / 182, 28 MP (85 MB) 227 /
// 229, 33 MP (98 MB) 262 / 262, 38 MP (113 MB) 294 / 287, 48 MP (146 MB) 355 / 321.
#ifndef JB2_FILL_IN_WRITE_BYTES
#define JB2_FILL_IN_WRITE_BYTES (100ull << 20)
#endif
constexpr size_t kFillInWriteBytes = JB2_FILL_IN_WRITE_BYTES;

struct ScanLayout {
    // upload block (offsets inside the staging block == inside the scratch buffer); the look-back
    // descriptors and the ticket of destuff travel as zeros with it
    size_t off_lut = 0, off_slow = 0, off_desc = 0;
    uint32_t desc_words = 0; // 32-bit words at off_desc
    // uploaded entropy-coded bytes
    size_t off_bytes = 0;
    // work area (shared by all scans of an image; scans run back to back on one stream)
    size_t off_ctrl = 0, off_destuffed = 0, off_coef = 0, off_dc = 0;
    size_t off_entries = 0, off_qpos = 0, off_seg_idx = 0, off_segs = 0, off_info = 0;
    size_t coef_bytes = 0, dc_bytes = 0;
    uint32_t num_chunks = 0, dc_tiles = 0, flag_words = 0, write_ctas = 0;
    uint32_t destuffed_bytes = 0;
    uint32_t num_dus = 0;
};

// Decode LUTs of one table set, keyed by the raw table specifications: expanding a DHT costs tens
// of microseconds, and consecutive images nearly always carry the same tables (same encoder).
struct LutSet {
    std::vector<uint8_t> key;
    std::vector<uint32_t> pool;
    HuffSlow slow[kMaxTables];
    uint32_t table_off[kMaxTables];
    uint32_t compact_words = 0;
};

inline void lut_key_of(const ScanPlan& sp, std::vector<uint8_t>& key)
{
    key.clear();
    for (int t = 0; t < sp.geom.num_tables; ++t) {
        const HuffSpec& h = sp.tables[t];
        key.push_back(h.is_dc ? 1 : 0);
        key.insert(key.end(), h.counts, h.counts + 16);
        key.insert(key.end(), h.vals, h.vals + 256);
    }
}

inline void build_lut_set(const ScanPlan& sp, const std::vector<uint8_t>& key, LutSet& ls)
{
    ls.pool.clear();
    for (int t = 0; t < sp.geom.num_tables; ++t) ls.table_off[t] = build_lut(sp.tables[t], ls.pool, ls.slow[t]);
    while (ls.pool.size() & 3) ls.pool.push_back(0); // staged with 16-byte loads
    ls.compact_words = 0; // what huff_write stages: header + one primary LUT + secondary
    for (int t = 0; t < sp.geom.num_tables; ++t)
        ls.compact_words += (uint32_t)(kLutHeader + kLut1Size) + ((ls.pool[ls.table_off[t] - kLutHeader + 1] + 3u) & ~3u);
    ls.key = key;
}

// false: more than 160 KiB of LUTs (does not fit the shared memory of the Huffman kernels)
inline bool apply_lut_set(const LutSet& ls, ScanGeom& g)
{
    for (int t = 0; t < g.num_tables; ++t) g.table_off[t] = ls.table_off[t];
    g.lut_pool_words = (uint32_t)ls.pool.size();
    g.lut_compact_words = ls.compact_words;
    return g.lut_pool_words <= 40 * 1024;
}

inline uint32_t destuff_desc_words(const ScanGeom& g)
{
    // (+ 15: a batched upload may start up to 15 bytes before the scan, kernels.hpp scan_lead)
    const size_t tiles = ((size_t)g.scan_avail + 15 + kDestuffTile - 1) / kDestuffTile;
    return (uint32_t)(4 * tiles + 4); // two 64-bit words per tile + the ticket
}

// Counts that follow from the number of subsequences / MCUs of a scan (no offsets).
inline void derive_counts(const ScanGeom& g, ScanLayout& L)
{
    L.num_dus = (uint32_t)g.total_mcus * (uint32_t)g.dus_per_mcu;
    L.num_chunks = (g.num_subs + kChunkSubs - 1) / kChunkSubs;
    L.dc_tiles = ((uint32_t)g.total_mcus + kDcTile - 1) / kDcTile;
    L.flag_words = (L.num_chunks + 1u) & ~1u; // keeps the 64-bit words behind it aligned
    L.write_ctas = g.num_subs / (32u / kWriteParts) + 1u; // huff_write runs >= one warp of subsequences per CTA
}

// Lays the work area of a scan out from `begin`; returns its end.
inline size_t plan_work_area(const ScanGeom& g, size_t begin, ScanLayout& L)
{
    size_t o = begin;
    derive_counts(g, L);
    L.off_ctrl = o; // control block, sync mailboxes and DC flags are cleared as one range
    o += align256(sizeof(Control) + ((size_t)L.flag_words + 4u * L.dc_tiles + 2u * L.write_ctas) * 4);
    L.off_destuffed = o;
    o += align256((size_t)g.num_subs * kSubBytes + 64);
    L.off_coef = o;
    L.coef_bytes = (size_t)L.num_dus * 128;
    o += align256(L.coef_bytes);
    L.off_dc = o;
    L.dc_bytes = ((size_t)L.num_dus * 2 + 15) & ~(size_t)15;
    o += align256(L.dc_bytes);
    L.off_entries = o;
    o += align256((size_t)g.num_subs * sizeof(SubEntry));
    L.off_qpos = o;
    o += align256(((size_t)g.num_subs + 1) * 4);
    L.off_seg_idx = o;
    o += align256((size_t)g.num_subs * 4);
    L.off_segs = o;
    o += align256(sizeof(SegDesc) * (size_t)g.num_segments);
    L.off_info = o;
    o += align256(sizeof(ScanInfo));
    return o;
}

// Device pointers of a scan whose regions sit at the layout's offsets from `base`.
inline ScanBufs make_bufs(const ScanLayout& L, void* base_ptr)
{
    uint8_t* base = static_cast<uint8_t*>(base_ptr);
    ScanBufs b;
    b.scan_bytes = base + L.off_bytes;
    b.scan_lead = 0;
    b.segs = reinterpret_cast<SegDesc*>(base + L.off_segs);
    b.info = reinterpret_cast<ScanInfo*>(base + L.off_info);
    b.destuff_desc = reinterpret_cast<unsigned long long*>(base + L.off_desc);
    b.destuff_words = L.desc_words;
    b.destuff_ticket = reinterpret_cast<uint32_t*>(base + L.off_desc) + (L.desc_words - 4);
    b.lut_pool = reinterpret_cast<const uint32_t*>(base + L.off_lut);
    b.slow = reinterpret_cast<const HuffSlow*>(base + L.off_slow);
    b.ctrl = reinterpret_cast<Control*>(base + L.off_ctrl);
    b.chunk_state = reinterpret_cast<uint32_t*>(base + L.off_ctrl + sizeof(Control));
    b.num_chunks = L.num_chunks;
    b.dc_agg = reinterpret_cast<unsigned long long*>(b.chunk_state + L.flag_words);
    b.write_agg = b.dc_agg + 2u * (size_t)L.dc_tiles;
    b.clear_words = (uint32_t)(sizeof(Control) / 4) + L.flag_words + 4u * L.dc_tiles + 2u * L.write_ctas;
    b.destuffed = reinterpret_cast<uint32_t*>(base + L.off_destuffed);
    b.coef = reinterpret_cast<int16_t*>(base + L.off_coef);
    b.coef_bytes = L.coef_bytes;
    b.fill_in_write = L.coef_bytes > kFillInWriteBytes ? 1u : 0u;
    b.dc = reinterpret_cast<int16_t*>(base + L.off_dc);
    b.dc_bytes = L.dc_bytes;
    b.entries = reinterpret_cast<SubEntry*>(base + L.off_entries);
    b.qpos = reinterpret_cast<uint32_t*>(base + L.off_qpos);
    b.seg_idx = reinterpret_cast<uint32_t*>(base + L.off_seg_idx);
    return b;
}

} // namespace jb2
#endif
