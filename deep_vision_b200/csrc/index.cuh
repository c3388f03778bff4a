// index.cuh — the resident chunk index handle shared by index.cu (build / query) and index_io.cu (sidecar save / load).
#pragma once
#include "common.cuh"
#include <vector>

// One segment = a contiguous range of chunks with its own sorted postings (<= IX_SEG_BYTES of text, so that the build's
// transient sort buffers stay bounded and a 180 GB shard is indexed piece by piece).  Passed BY VALUE to k_ix_score.
struct IxSeg {
    int64_t chunk_base = 0;             // position of the segment's first chunk inside the index
    int64_t n_chunks = 0, n_bytes = 0, n_post = 0, n_keys = 0;
    uint16_t* sym = nullptr;            // symbols of chunk c at sym[sym_off[c] .. + sym_cnt[c])
    int64_t* sym_off = nullptr;         // [n_chunks + 1] (= byte offsets rebased to 0)
    int32_t* sym_cnt = nullptr;         // [n_chunks]
    uint32_t* pc = nullptr;             // [n_post] chunk (segment-local) of every posting, sorted by (key, chunk, position)
    uint32_t* pv = nullptr;             // [n_post] position | symbol after the pair << 16
    uint32_t* ukeys = nullptr;          // [n_keys] distinct pair keys ascending
    int64_t* uoff = nullptr;            // [n_keys + 1]
};
struct dvb_index {
    int device = 0;
    int64_t n_chunks = 0, n_bytes = 0, n_post = 0, n_keys = 0;     // totals over the segments (n_keys: sum of per-segment distinct keys)
    std::vector<IxSeg> segs;
};
#define IX_SEG_BYTES (2000LL << 20)     // text bytes per segment: postings <= characters <= bytes < 2^31

void ix_free_seg(IxSeg& sg);
