// The fixed decomposition of the interval index space [0, n).
//
// The reference splits [a,b] by MPI rank (riemann.cpp:71-74, 4main.c:76-78) or
// by CUDA thread (cintegrate.cu:54-55), so its answer changes with the worker
// count.  Here the summation tree depends on n only:
//
//   point  -> (tile, row, column)     tile = rows x BQ_WIDTH points
//   tile   -> part                    BQ_NPART parts, near-equal runs of tiles
//   part   -> chunk                   2 parts per chunk (one per CTA of a cluster)
//   chunk  -> segment                 BQ_NCHUNK/8 chunks per segment
//   segment-> result                  8 segments, folded in order
//
// A rank of a world of 1/2/4/8 owns 8/world consecutive segments, a cluster
// owns whole chunks, a CTA whole parts, so any grid / world size produces the
// same bits.  Host and device both include this file.
#pragma once
#include <stdint.h>

#include "../include/b200quad.h"

#if defined(__CUDACC__)
#define BQ_HDI __host__ __device__ __forceinline__
#else
#define BQ_HDI inline
#endif

namespace bq {

constexpr int kThreads = 256;                 // threads per CTA (fixed: part of the tree)
constexpr int kCols = 8;                      // columns per thread
constexpr int kWidth = kThreads * kCols;      // points per row
constexpr int kMaxRows = 32;                  // rows per tile upper bound (one per lane)
constexpr int kFlushTiles = 4;                 // tiles a thread sums before a compensated flush
constexpr int kPartsPerChunk = 2;             // == cluster size when clusters are on
constexpr int kChunks = 1184;                 // 8 segments x 148 chunks: one per cluster on 8 GPUs
constexpr int kParts = kChunks * kPartsPerChunk;
constexpr int kSegments = BQ_SEGMENTS;
constexpr int kChunksPerSegment = kChunks / kSegments;
constexpr int kMaxTable = 4096;               // PROFILE table entries (smem)
constexpr uint64_t kMaxN = 1ull << 40;

enum PlanFlags : uint32_t {
    kFlagLocalOnly = 1u,     // skip the cross-GPU stage, leave segment pairs in seg_out
    kFlagSkipPublish = 2u,   // fault injection: never publish (peers time out)
    kFlagFastTrig = 4u,      // DIRECT: |angle| <= kFastTrigMax over [a, b], bounded sin/cos allowed
    kFlagRowK = 8u,          // ... and a row of a thread spans < 2^-10 rad: its samples share one multiple of pi/2
    kFlagFineGrid = 16u,     // ROTATE fp32: the 2048 columns of a row span < 2^-8 rad (TileAcc<float>: the C*Q sums stay tiny)
};

struct Plan {
    uint64_t n;
    double a, b, h, off;     // x_i = a + (i + off) * h
    double p[4];             // resolved integrand parameters
    double omega;            // angle = omega * x for trig integrands
    int32_t integrand, rule, dtype, eval;
    uint32_t rows;           // rows per full tile
    uint32_t table_len;      // PROFILE entries
    uint64_t tile_pts;
    uint64_t n_tiles;
    int32_t rank, world;
    uint32_t chunk_lo, chunk_hi;   // chunks owned by this rank
    uint32_t seg_lo, seg_hi;       // segments owned by this rank
    uint32_t flags;
    uint32_t timeout_ms;
};

BQ_HDI void part_tiles(const Plan& P, uint32_t part, uint64_t& lo, uint64_t& hi) {
    // n_tiles < 2^31 and part < 2^13, so the products fit in 64 bits.
    lo = (uint64_t(part) * P.n_tiles) / uint64_t(kParts);
    hi = (uint64_t(part + 1) * P.n_tiles) / uint64_t(kParts);
}

// Reference defaults: riemann.cpp:7-9.
constexpr double kTscale = 286.4788975;
constexpr double kAscale = 0.2365890;
constexpr double kVscale = 67.7777777;

// ROTATE, fused fp64 loop: (cos, sin) of the angle of row r of a flush group relative to the group's
// first tile base, angle(r * kWidth * h), r < kFlushTiles * kMaxRows.  They are launch constants, so
// they ride in the kernel parameters (constant bank) and enter the FMAs as uniform-register operands:
// one register read fewer per DFMA, which lifts the two-FMA stream from 86 % to 94 % of the fp64 issue
// rate (tools/dfma_bench.cu).  The host computes them with its libm for the GPU and the CPU backend
// alike (cached per thread for the last (integrand, h, tscale)).
struct RowConsts {
    double cs[kFlushTiles * kMaxRows][2];
};
void fill_row_consts(const Plan& P, RowConsts* out);

// Fills every derived field.  Returns a bq_status.
int make_plan(const bq_job& job, int rank, int world, uint32_t rows_override, int table_len,
              Plan* out);

}  // namespace bq
