// Plan of the fused cumulative-integration (scan) kernel; host + device.
//
// Capability of 4main.c:76-225 (fill, running sum, running sum of the running
// sum, cross-rank carry, broadcast) as a reduce-then-scan over fixed tiles:
//
//   tile    = kScanTile consecutive samples
//   segment = one of 8 near-equal runs of tiles (fixed by n alone)
//   rank r of a world of 1/2/4/8 owns 8/world consecutive segments
//
// Tile prefixes are built per segment with a fixed-shape block scan and the
// segment bases by an in-order fold over the 8 segment totals, so the tables
// are bit-identical for every world size.
#pragma once
#include <stdint.h>

#include "../include/b200quad.h"
#include "plan.h"

namespace bq {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;   // 2048 samples
constexpr uint64_t kScanMaxN = 1ull << 34;

enum ScanFlags : uint32_t {
    kScanReplicate = 1u,
    kScanSamples = 2u,
    kScanSkipPublish = 4u,
};

struct ScanPlan {
    uint64_t n;
    double x0, dx;
    double p[4];
    int32_t integrand;
    uint32_t table_len;
    int32_t rank, world;
    uint32_t flags;
    uint32_t timeout_ms;
    uint64_t n_tiles;
    uint64_t seg_tile[kSegments + 1];   // segment s = tiles [seg_tile[s], seg_tile[s+1])
    uint32_t seg_lo, seg_hi;            // segments owned by this rank
    uint64_t tile_lo, tile_hi;          // == seg_tile[seg_lo], seg_tile[seg_hi]
};

int make_scan_plan(const bq_scan_job& job, int rank, int world, int table_len, ScanPlan* out);

}  // namespace bq
