// Plan of the fused cumulative-integration (scan) kernel; host + device.
//
// Capability of 4main.c:76-225 (fill, running sum, running sum of the running
// sum, cross-rank carry, broadcast) as a reduce-then-scan over fixed tiles:
//
//   tile    = kScanTile consecutive samples
//   segment = one of 8 near-equal runs of tiles (fixed by n alone)
//   rank r of a world of 1/2/4/8 owns 8/world consecutive segments
//
// Inside a segment the tiles form groups of 32 and the groups supergroups of 32.
// The exclusive prefix of a tile is a fixed-shape expression of tile aggregates:
//
//   prefix(tile) = segbase(s) (+) [ (SGP(sg-1) (+) GTscan(sg)[gi-1]) (+) Ascan(g)[k-1] ]
//
//   Ascan(g)   Kogge-Stone inclusive scan over the (up to) 32 tile aggregates of group g
//   GT(g)      group total = last valid element of Ascan(g)
//   GTscan(sg) the same scan over the group totals of supergroup sg
//   SGP(sg)    in-order fold of the supergroup totals 0..sg of the segment
//   segbase(s) in-order fold of the totals of segments 0..s-1 (all ranks)
//
// Every element depends on earlier tiles only and on nothing but n, so (a) the
// tables are bit-identical for every world size, grid and schedule, and (b) a
// tile can form its prefix by itself as soon as its few inputs exist — which
// is what lets one GPU run the scan in a single pass (scan_kernel.cu).
#pragma once
#include <stdint.h>

#include "../include/b200quad.h"
#include "plan.h"

namespace bq {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;   // 2048 samples
constexpr uint64_t kScanMaxN = 1ull << 34;
constexpr int kScanGroup = 32;                          // tiles per group (one warp lane each)
#ifndef BQ_SCAN_SUPER
#define BQ_SCAN_SUPER 32
#endif
constexpr int kScanSuper = BQ_SCAN_SUPER;               // groups per supergroup (<= 32: one warp lane each)

enum ScanFlags : uint32_t {
    kScanReplicate = 1u,
    kScanSamples = 2u,
    kScanSkipPublish = 4u,
    kScanFastTrig = 8u,       // |angle| <= kFastTrigMax over the whole grid: bounded sin/cos (integrands.cuh)
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
    uint64_t seg_group0[kSegments + 1];   // groups of all segments before s (index of s's first group)
    uint64_t seg_super0[kSegments + 1];   // same for supergroups
    uint32_t seg_lo, seg_hi;            // segments owned by this rank
    uint64_t tile_lo, tile_hi;          // == seg_tile[seg_lo], seg_tile[seg_hi]
};

int make_scan_plan(const bq_scan_job& job, int rank, int world, int table_len, ScanPlan* out);

}  // namespace bq
