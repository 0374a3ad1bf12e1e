// The fused definite-integral kernel: ONE launch per GPU per integrate() call.
//
//   evaluate f on this rank's share of the n sample points        (hot loop,
//      no CTA-wide or cluster-wide synchronisation inside it)
//   -> per-thread dd totals -> warp shuffle -> per-warp part totals in smem
//   -> once per CTA: cross-warp combine, then the thread-block cluster pair
//      exchanges its part sums through DSMEM                       (chunk sums)
//   -> ticket; the last cluster folds chunk sums into segment sums
//   -> publish segment pairs into every peer's comm block over NVLink,
//      in-kernel epoch barrier, fixed-order fold, *h, rule epilogue -> result
//
// Replaces, in the reference: the per-thread loops + host sum of
// cintegrate.cu:47-98,135-138; riemann_sum + MPI_Send/Recv gather of
// riemann.cpp:29-44,76,82-85; MPI_Reduce/MPI_Barrier of 4main.c:134,137,200.
// No tensor-core work exists in this workload (no matrix product), so
// tcgen05/TMEM are not used; the one bulk transfer (the PROFILE table) goes
// through a 1-D TMA copy (cp.async.bulk -> UBLKCP).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdio.h>

#include <atomic>
#include <type_traits>

#include "device_state.h"
#include "ptx.cuh"
#include "quad_core.cuh"
#include "quad_launch.h"

// Rows of the ROTATE loop unrolled together.  Measured on B200 at n = 2^32, G intervals/s:
//   3-instruction loop (fp64 / fp32):  2: 5458 / 8281, 4: 5528 / 8660, 8: 5645 / 8684, 16: 5702 / 8616
//   fused 2-DFMA loop (fp64, SinF):    8: 7274, 16: 7374, 32: 7429; a second accumulator set per
//   column (16 / 8 rows): 7200 / 7079.  Pinning the column constants to aligned register quads
//   (so that constant and accumulator of a DFMA sit in different register banks) changed
//   nothing (7421): the loop is not limited by register-bank conflicts.  What it WAS limited by is
//   the number of register operands per DFMA: with the row constants as uniform-register operands
//   (PartWorker::run_uniform, fused fp64) the same two FMAs per sample run at 8 625.  BQ_ROT_UNROLL
//   now only shapes the general path (row bases from shared memory).
#ifndef BQ_ROT_UNROLL
#define BQ_ROT_UNROLL 16
#endif
#ifndef BQ_ROT_UNROLL_F32
#define BQ_ROT_UNROLL_F32 32    // = kPromoteRows: the promotion lands at the end of an unrolled body
#endif
#ifndef BQ_DIRECT_UNROLL
#define BQ_DIRECT_UNROLL 2      // rows of the DIRECT loop unrolled together (8 independent samples each)
#endif
#ifndef BQ_MIN_CTAS
#define BQ_MIN_CTAS 2
#endif

namespace bq {

template <class R>
struct RotUnroll {
    static constexpr int value = sizeof(R) == 8 ? BQ_ROT_UNROLL : BQ_ROT_UNROLL_F32;
};

template <class R>
struct Vec2;
template <>
struct __align__(16) Vec2<double> {
    double x, y;
};
template <>
struct __align__(8) Vec2<float> {
    float x, y;
};

__device__ __forceinline__ dd shfl_xor_dd(dd v, int off) {
    return dd{__shfl_xor_sync(0xffffffffu, v.hi, off), __shfl_xor_sync(0xffffffffu, v.lo, off)};
}

// Butterfly over the 32 lanes; every lane ends with the same pair.
__device__ __forceinline__ dd warp_reduce_dd(dd v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v = dd_add(v, shfl_xor_dd(v, off));
    return v;
}
// Same shape with un-normalised pairs (shorter dependent chain); see bq_dd.cuh.
__device__ __forceinline__ dd warp_reduce_lazy(dd v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v = dd_add_lazy(v, shfl_xor_dd(v, off));
    return v;
}

// (P_j - float P_j, Q_j - float Q_j) of column j of thread tid, derived afresh: for the ragged last row
// of the fused fp32 path (PartWorker::tile_rotate), once per launch.  Not inlined, so that its two
// libdevice sincos stay out of the register budget of the kernel's loops.
template <class F, class R>
__device__ __noinline__ double2 column_consts_lo(F f, double h, unsigned tid, int j, double cjv, double sjv) {
    SinCos mine, blk;
    sincos(f.angle(double(tid) * h), &mine.s, &mine.c);
    sincos(f.angle(double(j * kThreads) * h), &blk.s, &blk.c);
    double Pc, Qc;
    fused_consts(f, fused_norm<F, R>(f), rotate_sc(mine, blk), Pc, Qc);
    return make_double2(Pc - cjv, Qc - sjv);
}

template <class F, class R, int MODE>
struct PartWorker {
    static constexpr bool kRotate = (MODE == BQ_EVAL_ROTATE);
    static constexpr bool kFused = F::kFusedSum;   // see SinF / fused_consts
    // fused fp64: row constants as uniform operands, per-column values re-derived once per flush group
    static constexpr bool kUniformRows = kRotate && kFused && sizeof(R) == 8;

    const Plan& P;        // kernel parameter (constant bank)
    const F& f;
    Vec2<R>* wSC;         // this warp's private row bases: [2][kBufStride]
    double* wX0;          // this warp's private row x0:    [2][kBufStride] (integrands that need x)
    const double h, a;    // register copies for the hot loops
    const uint32_t rows;  // rows of every tile but the last
    const uint64_t last_tile, tile_pts;
    // column constants
    R cj[kCols], sj[kCols];
    const RowConsts* rowc; // fused fp64: the row constants in the kernel parameters (kUniformRows), else null
    const double* owed4;  // fused fp32: the CTA's four constants of the part's closing term, in shared memory:
                          // sums over all its columns of float(P_j), float(Q_j), P_j - float(P_j), Q_j - float(Q_j)
    const bool fine;      // kFlagFineGrid
    double dx[kCols];     // ROTATE: x offset of the column; DIRECT: column index as double
    SinCos rowrot;        // ROTATE: rotation of row `lane` relative to the tile base

    __device__ __forceinline__ PartWorker(const Plan& P_, const F& f_, Vec2<R>* sc, double* x0, double* owed,
                                          const RowConsts* rc)
        : P(P_), f(f_), wSC(sc), wX0(x0), h(P_.h), a(P_.a), rows(P_.rows),
          last_tile(P_.n_tiles - 1), tile_pts(P_.tile_pts), rowc(rc), owed4(owed),
          fine((P_.flags & kFlagFineGrid) != 0u) {
        const int lane = threadIdx.x & 31;
        double psum = 0.0, qsum = 0.0, plo = 0.0, qlo = 0.0;
        if (kRotate) {
            // Two sincos per thread for all kCols columns: the thread's own
            // offset, and (lane & 7)'s block offset shared through shuffles.
            SinCos mine, blk;
            sincos(f.angle(double(threadIdx.x) * h), &mine.s, &mine.c);
            sincos(f.angle(double((lane & (kCols - 1)) * kThreads) * h), &blk.s, &blk.c);
#pragma unroll
            for (int j = 0; j < kCols; ++j) {
                SinCos bj;
                bj.s = __shfl_sync(0xffffffffu, blk.s, j);
                bj.c = __shfl_sync(0xffffffffu, blk.c, j);
                const SinCos cs = rotate_sc(mine, bj);
                if constexpr (kFused) {
                    double Pc, Qc;
                    fused_consts(f, fused_norm<F, R>(f), cs, Pc, Qc);
                    cj[j] = fused_hi<R>(Pc, fine);
                    sj[j] = R(Qc);
                    const double ph = double(cj[j]), qh = double(sj[j]);
                    psum += ph, qsum += qh;
                    plo += Pc - ph, qlo += Qc - qh;
                } else {
                    cj[j] = R(cs.c);
                    sj[j] = R(cs.s);
                }
                dx[j] = double(threadIdx.x + kThreads * j) * h;
            }
            if constexpr (kUniformRows) {
                // row r of a tile is row r of a flush group: the host's constant (the general path only)
                rowrot = SinCos{rowc->cs[lane][1], rowc->cs[lane][0]};
            } else {
                sincos(f.angle(double(lane) * double(kWidth) * h), &rowrot.s, &rowrot.c);
            }
            if constexpr (RowDiffusion<R, kFused>::kOn) {
                // the four sums over the CTA's 2048 columns, once per launch: warp butterflies, then the
                // 8 warp sums in order (cpu_backend.cpp does the same); owed[0..3] the totals, owed[4..] scratch
#pragma unroll
                for (int off = 16; off >= 1; off >>= 1) {
                    psum += __shfl_xor_sync(0xffffffffu, psum, off);
                    qsum += __shfl_xor_sync(0xffffffffu, qsum, off);
                    plo += __shfl_xor_sync(0xffffffffu, plo, off);
                    qlo += __shfl_xor_sync(0xffffffffu, qlo, off);
                }
                const int warp = threadIdx.x >> 5;
                if (lane == 0) {
                    owed[4 + warp * 4 + 0] = psum, owed[4 + warp * 4 + 1] = qsum;
                    owed[4 + warp * 4 + 2] = plo, owed[4 + warp * 4 + 3] = qlo;
                }
                __syncthreads();
                if (threadIdx.x < 4) {
                    double t = 0.0;
                    for (int w = 0; w < kThreads / 32; ++w) t += owed[4 + w * 4 + threadIdx.x];
                    owed[threadIdx.x] = t;
                }
                __syncthreads();
            }
        } else {
#pragma unroll
            for (int j = 0; j < kCols; ++j) {
                dx[j] = double(threadIdx.x + kThreads * j);
                cj[j] = R(0);
                sj[j] = R(0);
            }
            rowrot = SinCos{0.0, 1.0};
        }
    }

    // Accurate sincos of the base angle of tile `first + lane` (one pass serves 32 tiles).
    __device__ __forceinline__ SinCos group_bases(uint64_t first, uint64_t hi) const {
        const uint64_t tile = first + (threadIdx.x & 31);
        SinCos g{0.0, 1.0};
        if (tile < hi) {
            const double xt = fma(double(tile * tile_pts) + P.off, h, a);
            sincos(f.angle(xt), &g.s, &g.c);
        }
        return g;
    }

    // Rows of a flush group: kFlushTiles tiles x kMaxRows rows.  The row bases of a
    // whole group sit contiguously in this warp's buffer, so a group of full
    // tiles is one software-pipelined loop of kGroupRows rows.
    static constexpr int kGroupRows = kFlushTiles * kMaxRows;
    static constexpr int kBufStride = kGroupRows + 2;   // +1 prefetch slot, kept even

    // Row bases of the group of `nt` tiles starting at tile t0: for each tile,
    // lane r rotates the tile base by r rows (4 fp64 ops) and parks the pair in
    // this warp's private buffer.  `grp` carries the accurate tile bases of the
    // current run of 32 tiles (refreshed here when t0 starts a new run).
    __device__ __forceinline__ void group_row_bases(uint32_t t0, uint32_t nt, uint32_t lo, uint32_t hi,
                                                    SinCos& grp, int buf, RowDiffusion<R, kFused>& diff) const {
        const int lane = threadIdx.x & 31;
        if (((t0 - lo) & 31u) == 0u) grp = group_bases(t0, hi);
        Vec2<R>* sc = wSC + buf * kBufStride + lane;
        double* x0 = wX0 + buf * kBufStride + lane;
#pragma unroll
        for (uint32_t k = 0; k < uint32_t(kFlushTiles); ++k) {
            if (k < nt) {
                const int src = int((t0 + k - lo) & 31u);
                const SinCos tb{__shfl_sync(0xffffffffu, grp.s, src), __shfl_sync(0xffffffffu, grp.c, src)};
                SinCos rb = rotate_sc(tb, rowrot);
                // fused fp32: the row values of whole tiles go on a grid, this lane's row stream carrying
                // what it owes from tile to tile (quad_core.cuh: RowDiffusion); a plain rounding otherwise
                R rs, rc;
                if (kFused && fused_norm<F, R>(f).swap) rb = SinCos{rb.c, rb.s};   // fp32, cos-type integrand: streams (C, S)
                uint32_t full = rows;      // full rows of the tile: all of them, but for the grid's last tile
                if (RowDiffusion<R, kFused>::kOn && uint64_t(t0 + k) == last_tile) full = tile_geom(P, last_tile).full_rows;
                diff.round(rb, uint32_t(lane) < full, rs, rc);
                sc[k * kMaxRows] = Vec2<R>{rs, rc};
                if (F::kNeedX)
                    x0[k * kMaxRows] =
                        fma(double(uint64_t(t0 + k) * tile_pts + uint64_t(lane) * kWidth) + P.off, h, a);
            }
        }
    }

    // `nrows` full rows starting at row-base slot 0 of sc_row / x0_row.
    // Software-pipelined: the next row's base is in flight while this row's
    // 24 fp64 instructions issue (slot nrows is valid memory).
    template <class Count>
    __device__ __forceinline__ void rows_rotate(const Vec2<R>* sc_row, const double* x0_row, Count nrows,
                                                TileAcc<R>& acc, Vec2<R>& sc, double& x0) const {
        sc = sc_row[0];
        x0 = F::kNeedX ? x0_row[0] : 0.0;
        constexpr int kUnroll = RotUnroll<R>::value;
#pragma unroll kUnroll
        for (Count r = 0; r < nrows; ++r) {
            const Vec2<R> sc_next = sc_row[r + 1];
            const double x0_next = F::kNeedX ? x0_row[r + 1] : 0.0;
            if constexpr (kFused) {
                acc.add_rot_row(sc.x, sc.y, cj, sj);
                if ((r & (kPromoteRows - 1)) == kPromoteRows - 1) acc.promote(fine);   // fp32 only; no-op in fp64
            } else {
                R v[kCols];
#pragma unroll
                for (int j = 0; j < kCols; ++j)
                    v[j] = point_rotate<F, R>(f, sc.x, sc.y, cj[j], sj[j], x0, dx[j]);
                acc.add_row(v);
            }
            sc = sc_next;
            x0 = x0_next;
        }
    }

    // `nblk` blocks of kUnroll full rows starting at slot 0: the same pipeline with a body that is
    // always whole, for any number of full tiles in a flush group (no remainder loop).
    __device__ __forceinline__ void rows_rotate_blocks(const Vec2<R>* sc_row, const double* x0_row, int nblk,
                                                       TileAcc<R>& acc) const {
        constexpr int kUnroll = RotUnroll<R>::value;
        static_assert(kMaxRows % kUnroll == 0 && (std::is_same<R, double>::value || kUnroll % kPromoteRows == 0),
                      "a tile is whole blocks, an fp32 run ends with a block");
        Vec2<R> sc = sc_row[0];
        double x0 = F::kNeedX ? x0_row[0] : 0.0;
#pragma unroll 1
        for (int b = 0; b < nblk; ++b) {
#pragma unroll
            for (int r = 0; r < kUnroll; ++r) {
                const Vec2<R> sc_next = sc_row[b * kUnroll + r + 1];
                const double x0_next = F::kNeedX ? x0_row[b * kUnroll + r + 1] : 0.0;
                if constexpr (kFused) {
                    acc.add_rot_row(sc.x, sc.y, cj, sj);
                    if ((r & (kPromoteRows - 1)) == kPromoteRows - 1) acc.promote(fine);   // fp32 only
                } else {
                    R v[kCols];
#pragma unroll
                    for (int j = 0; j < kCols; ++j)
                        v[j] = point_rotate<F, R>(f, sc.x, sc.y, cj[j], sj[j], x0, dx[j]);
                    acc.add_row(v);
                }
                sc = sc_next;
                x0 = x0_next;
            }
        }
    }

    // ROTATE, any tile (short last tile, tiles of fewer than kMaxRows rows).
    __device__ __forceinline__ void tile_rotate(uint32_t tile, const Vec2<R>* sc_row, const double* x0_row,
                                                TileAcc<R>& acc) const {
        uint32_t full_rows = rows, rem = 0;
        if (tile == last_tile) {
            const TileGeom g = tile_geom(P, tile);
            full_rows = g.full_rows, rem = g.rem;
        }
        Vec2<R> sc;
        double x0;
        rows_rotate<uint32_t>(sc_row, x0_row, full_rows, acc, sc, x0);
        if (rem) {   // sc / x0 already hold row full_rows
#pragma unroll
            for (int j = 0; j < kCols; ++j)
                if (threadIdx.x + kThreads * j < rem) {
                    if constexpr (kFused) {
                        acc.add_rot(j, sc.x, sc.y, cj[j], sj[j]);
                        if constexpr (RowDiffusion<R, kFused>::kOn) {
                            // The ragged last row is outside the parts' closing terms (not every column
                            // uses it): what the fp32 rounding of its column constants leaves — on a fine
                            // grid P_j is taken as 1.0f, up to 8e-6 off — is added here, in fp64.  Once per
                            // launch, so the constants are simply derived again.
                            const double2 lo = column_consts_lo<F, R>(f, h, threadIdx.x, j, double(cj[j]), double(sj[j]));
                            acc.d[j] += fma(double(sc.x), lo.x, double(sc.y) * lo.y);
                        }
                    } else
                        acc.add_one(j, point_rotate<F, R>(f, sc.x, sc.y, cj[j], sj[j], x0, dx[j]));
                }
        }
    }

    // DIRECT: adds every sample of `tile` into `acc`.  Only the globally last tile can be short.
    template <int FAST>
    __device__ __forceinline__ void tile_direct(uint64_t tile, TileAcc<R>& acc) const {
        TileGeom g;
        if (tile == last_tile) {
            g = tile_geom(P, tile);
        } else {
            g.start = tile * tile_pts;
            g.full_rows = rows;
            g.rem = 0;
        }
        double rb = row_base_index(P, g, 0);
        constexpr int kDirectUnroll = BQ_DIRECT_UNROLL;
#pragma unroll kDirectUnroll
        for (uint32_t r = 0; r < g.full_rows; ++r) {
            R v[kCols];
            RowDirect<F, R, FAST>::eval(f, rb, dx, h, a, v);
            acc.add_row(v);
            rb += double(kWidth);
        }
        if (g.rem) {
#pragma unroll
            for (int j = 0; j < kCols; ++j)
                if (threadIdx.x + kThreads * j < g.rem)
                    acc.add_one(j, point_direct<F, R, FAST>(f, rb, dx[j], h, a));
        }
    }

    // Fused fp64 (sin, acc, vel): the samples of a flush group of full tiles, as
    //   f(row r, column j) = A_j * cos(rho_r) + B_j * sin(rho_r),   rho_r = angle(r * kWidth * h), r < 128,
    //   A_j = Sb * P_j + Cb * Q_j,  B_j = Cb * P_j - Sb * Q_j       (Sb, Cb): sincos of the group's base angle
    // i.e. the rotation is applied to the column side once per group (4 fp64 instructions per column per
    // 128 rows) and the row side is the same 128 constants for every group of the launch.  Those sit in
    // the kernel parameters, so the two FMAs of a sample read them as uniform-register operands
    // (SASS: LDCU.64 UR, c[0x0][UR+imm]; DFMA R, R, UR, R) — two register reads per DFMA instead of
    // three, no shared memory and no warp synchronisation in the loop.  The group that holds the
    // grid's last tile (which may be ragged) takes the general path.
    __device__ __forceinline__ dd run_uniform(uint32_t lo, uint32_t hi) const {
        dd total{0.0, 0.0};
        if (lo >= hi) return total;
        TileAcc<R> acc;
        acc.clear();
        const uint32_t last = uint32_t(last_tile);
        SinCos grp{0.0, 1.0};
        RowDiffusion<R, kFused> diff;   // (no state in fp64)
#pragma unroll 1
        for (uint32_t t = lo; t < hi; t += kFlushTiles) {
            const uint32_t nt = hi - t < uint32_t(kFlushTiles) ? hi - t : uint32_t(kFlushTiles);
            if (t + nt - 1 < last) {          // full tiles of `rows` rows each: nt * rows consecutive rows
                if (((t - lo) & 31u) == 0u) grp = group_bases(t, hi);
                const int src = int((t - lo) & 31u);
                const double sb = __shfl_sync(0xffffffffu, grp.s, src), cb = __shfl_sync(0xffffffffu, grp.c, src);
                R A[kCols], B[kCols];
#pragma unroll
                for (int j = 0; j < kCols; ++j) {
                    A[j] = fma(sb, cj[j], cb * sj[j]);
                    B[j] = fma(cb, cj[j], -(sb * sj[j]));
                }
                if (rows == uint32_t(kMaxRows)) {
#pragma unroll 1
                    for (uint32_t k = 0; k < nt; ++k) {
#pragma unroll
                        for (int r = 0; r < kMaxRows; ++r) {
                            const double cr = rowc->cs[k * kMaxRows + r][0], sr = rowc->cs[k * kMaxRows + r][1];
                            acc.add_rot_row(cr, sr, A, B);      // a_j = fma(sr, B_j, fma(cr, A_j, a_j))
                        }
                    }
                } else {                           // tiles of fewer than kMaxRows rows (n below ~1.5e8)
                    const uint32_t nq = nt * rows;
#pragma unroll 4
                    for (uint32_t q = 0; q < nq; ++q) acc.add_rot_row(rowc->cs[q][0], rowc->cs[q][1], A, B);
                }
            } else {
                group_row_bases(t, nt, lo, hi, grp, 0, diff);
                __syncwarp();
                for (uint32_t k = 0; k < nt; ++k) tile_rotate(t + k, wSC + k * kMaxRows, wX0 + k * kMaxRows, acc);
                __syncwarp();
            }
            total = dd_acc_lazy(total, acc.flush_fused(fine));
            acc.clear();
        }
        return total;
    }

    // All tiles of one part; returns this thread's dd total.  In ROTATE mode a
    // warp is self-sufficient (its own bases, its own shared-memory buffer), so
    // the tile loop has no CTA-wide barrier, and a flush group of full tiles
    // runs as one loop: one warp-level sync and one compensated flush per
    // kFlushTiles tiles.
    __device__ __forceinline__ dd run(uint64_t lo64, uint64_t hi64) const {   // tiles [lo64, hi64) of a part
        dd total{0.0, 0.0};
        TileAcc<R> acc;
        acc.clear();
        if constexpr (kUniformRows) {
            return run_uniform(uint32_t(lo64), uint32_t(hi64));
        } else if (kRotate) {
            const uint32_t lo = uint32_t(lo64), hi = uint32_t(hi64);   // n_tiles < 2^31
            if (lo >= hi) return total;
            const uint32_t last = uint32_t(last_tile);
            SinCos grp{0.0, 1.0};
            RowDiffusion<R, kFused> diff;
            int buf = 0;
            group_row_bases(lo, hi - lo < uint32_t(kFlushTiles) ? hi - lo : uint32_t(kFlushTiles), lo, hi, grp, 0, diff);
            __syncwarp();
            for (uint32_t t = lo; t < hi; t += kFlushTiles) {
                const uint32_t nt = hi - t < uint32_t(kFlushTiles) ? hi - t : uint32_t(kFlushTiles);
                if (t + kFlushTiles < hi) {
                    const uint32_t left = hi - t - kFlushTiles;
                    group_row_bases(t + kFlushTiles, left < uint32_t(kFlushTiles) ? left : uint32_t(kFlushTiles),
                                    lo, hi, grp, buf ^ 1, diff);
                }
                const Vec2<R>* sc_row = wSC + buf * kBufStride;
                const double* x0_row = wX0 + buf * kBufStride;
                if (rows == uint32_t(kMaxRows) && t + nt - 1 < last) {
                    // every tile of the group is full (also the 1-3 tiles that end a part): one
                    // software-pipelined loop over all its rows
                    rows_rotate_blocks(sc_row, x0_row, int(nt) * (kMaxRows / RotUnroll<R>::value), acc);
                } else {
                    for (uint32_t k = 0; k < nt; ++k)
                        tile_rotate(t + k, sc_row + k * kMaxRows, x0_row + k * kMaxRows, acc);
                }
                total = dd_acc_lazy(total, kFused ? fused_norm<F, R>(f).amp * acc.flush_fused(fine) : acc.flush());
                acc.clear();
                __syncwarp();
                buf ^= 1;
            }
            if constexpr (RowDiffusion<R, kFused>::kOn) {
                // What the row streams still owe the sums of this part.  Every warp carries the same 32
                // streams (they walk the same tiles), so warp 0 settles them: lane r adds its stream's
                // term for ALL the CTA's columns — no shuffle, four FMAs per part.
                if (threadIdx.x < 32)
                    total = dd_acc_lazy(total, fused_norm<F, R>(f).amp * diffusion_owed(diff.es, diff.ec, diff.ss, diff.sc,
                                                                                      owed4[0], owed4[1], owed4[2], owed4[3]));
            }
        } else {
            // uniform: kernel parameters
            const int fast = !(F::kTrig && (P.flags & kFlagFastTrig)) ? 0 : ((P.flags & kFlagRowK) ? 2 : 1);
            for (uint64_t t = lo64; t < hi64; ++t) {
                if (fast == 2)
                    tile_direct<2>(t, acc);
                else if (fast == 1)
                    tile_direct<1>(t, acc);
                else
                    tile_direct<0>(t, acc);
                if (((t - lo64) & (kFlushTiles - 1)) == kFlushTiles - 1 || t + 1 == hi64) {
                    total = dd_acc_lazy(total, acc.flush());
                    acc.clear();
                }
            }
        }
        return total;   // un-normalised; renormalised once per part after the cross-warp combine
    }
};

// Parts a CTA reduces in one go: its share of up to kBatchChunks chunks
// (one part per chunk when a cluster pair splits the chunk, else both).
constexpr int kBatchChunks = 8;
constexpr int kBatchSlots = kBatchChunks * kPartsPerChunk;

template <bool ROWC>
struct QuadParamsT {
    Plan P;            // 152 bytes, rides in the kernel parameter buffer
    DeviceState st;
    __host__ __device__ const RowConsts* rowc() const { return nullptr; }
};
template <>
struct QuadParamsT<true> {
    Plan P;
    DeviceState st;
    RowConsts rc;      // 2 KB more for the fused fp64 ROTATE kernels (PartWorker::run_uniform)
    __host__ __device__ const RowConsts* rowc() const { return &rc; }
};
template <class F, class R, int MODE>
using QuadParams = QuadParamsT<PartWorker<F, R, MODE>::kUniformRows>;

template <class F, class R, int MODE>
__global__ void __launch_bounds__(kThreads, BQ_MIN_CTAS) quad_kernel(const __grid_constant__ QuadParams<F, R, MODE> prm) {
    const Plan& P = prm.P;
    const DeviceState& st = prm.st;
    // warp-private row-base buffers (ROTATE only), sized by rotate_smem_bytes()
    extern __shared__ __align__(16) unsigned char sDyn[];
    using Worker = PartWorker<F, R, MODE>;
    Vec2<R>* const sSC = reinterpret_cast<Vec2<R>*>(sDyn);
    double* const sX0 = reinterpret_cast<double*>(sDyn + sizeof(Vec2<R>) * (kThreads / 32) * 2 * Worker::kBufStride);
    __shared__ dd sWarpTot[kBatchSlots][kThreads / 32];   // per-warp part totals, reduced once per batch
    __shared__ dd sMine[kBatchSlots];                      // this CTA's part sums of the batch
    __shared__ uint32_t sLo[kBatchSlots], sHi[kBatchSlots];   // tile range of each slot's part
    __shared__ __align__(16) dd sMail[kBatchChunks];       // DSMEM mailbox (leader CTA)
    __shared__ dd sSeg[kSegments];
    __shared__ __align__(8) bq_result sRes;                // result block, assembled by lane 0 of the last warp
    __shared__ int sIsLast;
    constexpr bool kIsProfile = std::is_same<F, ProfileF>::value;
    __shared__ __align__(16) double sTable[kIsProfile ? kMaxTable : 2];
    __shared__ __align__(8) uint64_t sBar;

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const uint32_t crank = ptx::cluster_ctarank();
    const uint32_t csize = ptx::cluster_nctarank();

    if (blockIdx.x == 0 && tid == 0) st.result->t_start_ns = ptx::globaltimer();
    uint64_t trace_t0 = 0, trace_tiles = 0;
    if (st.trace && tid == 0) trace_t0 = ptx::globaltimer();

    if (kIsProfile) {
        // One 1-D TMA bulk copy brings the whole table (14.4 KB by default).
        if (tid == 0) ptx::mbar_init(&sBar, 1);
        __syncthreads();
        const uint32_t bytes = ((P.table_len + 1u) & ~1u) * 8u;
        if (tid == 0) {
            ptx::mbar_expect_tx(&sBar, bytes);
            ptx::tma_bulk_g2s(sTable, st.table, bytes, &sBar);
        }
        ptx::mbar_wait(&sBar, 0);
    }

    const F f = make_functor<F>(P, sTable);
    __shared__ double sOwed[RowDiffusion<R, Worker::kFused>::kOn && Worker::kRotate ? 4 + 4 * (kThreads / 32) : 1];
    const Worker worker(P, f, sSC + warp * 2 * Worker::kBufStride,
                        sX0 + (F::kNeedX ? warp * 2 * Worker::kBufStride : 0), sOwed, prm.rowc());
    // absolute %globaltimer stamps of this CTA (kept by thread 0; only the finishing CTA's are used)
    uint64_t ts_prologue = 0, ts_tiles = 0;
    // The call number of the cross-GPU exchange is fetched now, by the thread that may need it at
    // the very end: the L2 miss (~0.7 us) hides behind the tile loop instead of sitting in the tail.
    uint64_t epoch_pre = 0;
    if (tid == 0) {
        ts_prologue = ptx::globaltimer();
        if (P.world > 1 && !(P.flags & kFlagLocalOnly)) epoch_pre = ptx::ld_relaxed_sys_u64(&st.comm->epoch);
    }

    const uint32_t cluster_id = blockIdx.x / csize;
    const uint32_t n_clusters = gridDim.x / csize;

    // ---- chunks ------------------------------------------------------------
    // A cluster owns a contiguous run of chunks (for small n the few chunks that
    // hold a tile are evenly spaced; a strided assignment would pile them onto a
    // handful of clusters).  With a cluster pair, CTA q takes part q of every
    // chunk.  Nothing in the tile loop synchronises beyond a warp: every warp
    // parks its total of a part in shared memory and moves on, and the
    // CTA-wide / cluster-wide combine happens once per batch of kBatchChunks
    // chunks — one __syncthreads and one cluster barrier per launch on the
    // default grid, instead of one of each per chunk.
    const uint32_t per_cluster = (P.chunk_hi - P.chunk_lo + n_clusters - 1) / n_clusters;
    const uint32_t my_lo = P.chunk_lo + cluster_id * per_cluster;
    const uint32_t my_hi = my_lo + per_cluster < P.chunk_hi ? my_lo + per_cluster : P.chunk_hi;
    const uint32_t pp = (csize == 2) ? 1u : uint32_t(kPartsPerChunk);   // parts per chunk done here
    for (uint32_t b_lo = my_lo; b_lo < my_hi; b_lo += kBatchChunks) {
        const uint32_t nb = my_hi - b_lo < uint32_t(kBatchChunks) ? my_hi - b_lo : uint32_t(kBatchChunks);
        // Tile ranges of the batch's parts: one thread per slot works them out (two 64-bit
        // divisions each) instead of every thread of the CTA repeating all of them — for small
        // n, where most parts are empty, that repetition was most of a CTA's life.
        if (uint32_t(tid) < nb * pp) {
            const uint32_t k = uint32_t(tid) / pp, q = uint32_t(tid) % pp;
            const uint32_t part = (b_lo + k) * kPartsPerChunk + (csize == 2 ? crank : q);
            uint64_t lo, hi;
            part_tiles(P, part, lo, hi);
            sLo[tid] = uint32_t(lo), sHi[tid] = uint32_t(hi);   // n_tiles < 2^31
        }
        __syncthreads();
        for (uint32_t slot = 0; slot < nb * pp; ++slot) {
            const uint32_t lo = sLo[slot], hi = sHi[slot];
            trace_tiles += hi - lo;
            dd t{0.0, 0.0};
            if (lo < hi) t = warp_reduce_lazy(worker.run(lo, hi));   // small n: most parts hold no tile
            if (lane == 0) sWarpTot[slot][warp] = t;
        }
        __syncthreads();
        // Part sums: warp w finishes slots w, w + 8 (the 8-lane butterfly + one
        // renormalisation).
        for (uint32_t s = warp; s < nb * pp; s += kThreads / 32) {
            dd r = lane < kThreads / 32 ? sWarpTot[s][lane] : dd{0.0, 0.0};
#pragma unroll
            for (int off = 4; off >= 1; off >>= 1) r = dd_add_lazy(r, shfl_xor_dd(r, off));
            r = dd_renorm(r);
            if (lane == 0) {
                if (csize == 2 && crank == 1)
                    ptx::st_dsmem_f64x2(ptx::mapa(&sMail[s], 0), r.hi, r.lo);   // partner's mailbox
                else
                    sMine[s] = r;
            }
        }
        if (csize == 2)
            ptx::cluster_sync_all();    // mailbox complete (release/acquire), also a CTA barrier
        else
            __syncthreads();
        if (crank == 0 && uint32_t(tid) < nb) {
            const dd p0 = sMine[tid * pp];
            const dd p1 = (csize == 2) ? sMail[tid] : sMine[tid * pp + 1];
            st.partials[b_lo + tid] = dd_add(p0, p1);
        }
        if (b_lo + kBatchChunks < my_hi) {   // another batch reuses the slots and the mailbox
            if (csize == 2)
                ptx::cluster_sync_all();
            else
                __syncthreads();
        }
    }
    if (tid == 0) ts_tiles = ptx::globaltimer();
    if (st.trace && tid == 0) {
        uint32_t smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        uint64_t* t = st.trace + 4 * blockIdx.x;
        t[0] = trace_t0, t[1] = ptx::globaltimer(), t[2] = smid, t[3] = trace_tiles;
    }
    if (crank != 0) return;

    // ---- ticket: who is last? ------------------------------------------------
    __syncthreads();                 // every chunk sum of this cluster is stored
    if (tid == 0) {
        __threadfence();
        const unsigned int t = atomicAdd(st.ticket, 1u);
        sIsLast = (t == n_clusters - 1);
    }
    __syncthreads();
    if (!sIsLast) return;
    __threadfence();
    uint64_t ts_ticket = 0, ts_fold = 0, ts_pub = 0, ts_bar = 0;
    if (tid == 0) ts_ticket = ptx::globaltimer();
    // kernel-entry stamp of CTA 0 (ordered before this point by the ticket chain): fetched now so
    // that the load is long done when the block is assembled
    const uint64_t t_start = *reinterpret_cast<volatile const uint64_t*>(&st.result->t_start_ns);

    // ---- segment sums: one warp per owned segment, fixed order ---------------
    if (tid < kSegments) sSeg[tid] = dd{0.0, 0.0};
    __syncthreads();
    {
        const uint32_t seg = P.seg_lo + warp;
        if (seg < P.seg_hi) {
            constexpr int kPerLane = (kChunksPerSegment + 31) / 32;
            const uint32_t c0 = seg * kChunksPerSegment;
            // all loads in flight first, then the same in-order fold as ever
            double hi[kPerLane], lo[kPerLane];
#pragma unroll
            for (int i = 0; i < kPerLane; ++i) {
                const uint32_t c = c0 + lane + 32 * i;
                hi[i] = lo[i] = 0.0;
                if (c < c0 + kChunksPerSegment)
                    ptx::ld_relaxed_sys_f64x2(reinterpret_cast<const double*>(st.partials + c), hi[i], lo[i]);
            }
            dd acc{0.0, 0.0};
#pragma unroll
            for (int i = 0; i < kPerLane; ++i)
                if (c0 + lane + 32 * i < c0 + kChunksPerSegment) acc = dd_add(acc, dd{hi[i], lo[i]});
            acc = warp_reduce_dd(acc);
            if (lane == 0) sSeg[seg] = acc;
        }
    }
    __syncthreads();
    if (warp != 0) return;
    if (lane == 0) ts_fold = ts_pub = ts_bar = ptx::globaltimer();

    int status = BQ_OK;
    uint32_t spins = 0;
    const bool exchange = P.world > 1 && !(P.flags & kFlagLocalOnly);
    if (exchange) {
        // ---- cross-GPU: every rank's segment pairs land in every rank's comm block as LL words
        // (payload | call number per 8-byte word): the stores ARE the message, nothing is fenced
        // or counted, and the wait below polls the 32 words of this call's parity directly.
        const uint64_t epoch = __shfl_sync(0xffffffffu, epoch_pre, 0);   // lane 0 of warp 0 is thread 0
        const int par = int(epoch & 1);
        const uint32_t flag = uint32_t(epoch % 0xffffffffull) + 1u;   // never 0 (0 = a cleared block)
        if (!(P.flags & kFlagSkipPublish)) {
            if (st.mc_comm) {
                // NVLS: one multimem.st per word, fanned out by the switch to all ranks (self included)
                if (uint32_t(lane) < P.seg_hi - P.seg_lo) {
                    const uint32_t s = P.seg_lo + lane;
                    ptx::ll_store_multicast(&st.mc_comm->ll[par][s][0], sSeg[s].hi, sSeg[s].lo, flag);
                }
            } else if (lane < P.world) {
                CommBlock* peer = st.peers[lane];
                for (uint32_t s = P.seg_lo; s < P.seg_hi; ++s)
                    ptx::ll_store(&peer->ll[par][s][0], sSeg[s].hi, sSeg[s].lo, flag);
            }
        }
        const uint64_t t0 = ptx::globaltimer();
        if (lane == 0) ts_pub = t0;
        const uint64_t limit = uint64_t(P.timeout_ms) * 1000000ull;
        static_assert(kSegments * 4 == 32, "one LL word per lane");
        const uint64_t* word = &st.comm->ll[par][0][0] + lane;
        uint64_t w;
        for (;;) {
            w = ptx::ld_relaxed_sys_u64(word);
            if (__all_sync(0xffffffffu, uint32_t(w >> 32) == flag)) break;
            if ((++spins & 63u) == 0u) {
                const bool late = ptx::globaltimer() - t0 > limit;
                if (__any_sync(0xffffffffu, late)) {
                    status = BQ_ERR_TIMEOUT;
                    break;
                }
            }
        }
        if (lane == 0) ts_bar = ptx::globaltimer();
        // lane s < 8 reassembles segment s from lanes 4s .. 4s+3
        const uint32_t v = uint32_t(w);
        const int src = (lane & (kSegments - 1)) * 4;
        const uint64_t h = uint64_t(__shfl_sync(0xffffffffu, v, src)) | (uint64_t(__shfl_sync(0xffffffffu, v, src + 1)) << 32);
        const uint64_t l = uint64_t(__shfl_sync(0xffffffffu, v, src + 2)) | (uint64_t(__shfl_sync(0xffffffffu, v, src + 3)) << 32);
        if (lane < kSegments) sSeg[lane] = dd{__longlong_as_double((long long)h), __longlong_as_double((long long)l)};
        __syncwarp();
        if (lane == 0) st.comm->epoch = epoch + 1;
    }

    // The result block is assembled in shared memory and leaves in two warp-wide bursts (device
    // block, host block): no thread reads back what another one just stored to global memory.
    if (lane == 0) {
        double hi[kSegments], lo[kSegments];
        for (int s = 0; s < kSegments; ++s) hi[s] = sSeg[s].hi, lo[s] = sSeg[s].lo;
        if (st.seg_out) {
            for (int s = 0; s < kSegments; ++s) st.seg_out[2 * s] = hi[s], st.seg_out[2 * s + 1] = lo[s];
        }
        finish_result<F>(P, f, hi, lo, &sRes);
        if (status != BQ_OK) sRes.value = nan("");
        if (st.value_out) *st.value_out = sRes.value;
        sRes.t_start_ns = t_start;
        sRes.status = status;
        sRes.device_ms = 0.f;
        sRes.seq = st.seq;
        sRes.reserved = 0;
        *st.ticket = 0u;
        const uint64_t t_end = ptx::globaltimer();
        sRes.t_end_ns = t_end;
        auto rel = [&](uint64_t t) { return t > t_start ? uint32_t(t - t_start) : 0u; };
        sRes.stamp_ns[BQ_STAMP_PROLOGUE] = rel(ts_prologue);
        sRes.stamp_ns[BQ_STAMP_TILES] = rel(ts_tiles);
        sRes.stamp_ns[BQ_STAMP_TICKET] = rel(ts_ticket);
        sRes.stamp_ns[BQ_STAMP_FOLD] = rel(ts_fold);
        sRes.stamp_ns[BQ_STAMP_PUBLISHED] = rel(ts_pub);
        sRes.stamp_ns[BQ_STAMP_BARRIER] = rel(ts_bar);
        sRes.stamp_ns[BQ_STAMP_RESULT] = rel(t_end);
        sRes.stamp_ns[BQ_STAMP_SPINS] = spins;
    }
    __syncwarp();
    static_assert(sizeof(bq_result) % 8 == 0 && sizeof(bq_result) / 8 <= 32, "result block = one word per lane");
    constexpr int kWords = int(sizeof(bq_result) / 8);
    constexpr int kSeqWord = int(offsetof(bq_result, seq) / 8);
    const unsigned long long word64 = lane < kWords ? reinterpret_cast<const unsigned long long*>(&sRes)[lane] : 0ull;
    if (lane < kWords) reinterpret_cast<unsigned long long*>(st.result)[lane] = word64;
    // The device -> host leg of the call: the block goes straight into the context's pinned,
    // mapped host block (a single PCIe write burst) — no copy engine operation follows the
    // kernel.  The call number goes last, after a system-scope fence, so a host that watches
    // that word sees a complete block.
    if (st.result_host) {
        volatile unsigned long long* dst = reinterpret_cast<volatile unsigned long long*>(st.result_host);
        if (lane < kWords && lane != kSeqWord) dst[lane] = word64;
        __threadfence_system();
        __syncwarp();
        if (lane == 0) dst[kSeqWord] = st.seq;
    }
}

// ---- dispatch ----------------------------------------------------------------

// Dynamic shared memory of one CTA: the warp-private row-base buffers of ROTATE.
template <class F, class R, int MODE>
constexpr size_t rotate_smem_bytes() {
    using W = PartWorker<F, R, MODE>;
    return MODE == BQ_EVAL_ROTATE
               ? size_t(kThreads / 32) * 2 * W::kBufStride * (sizeof(Vec2<R>) + (F::kNeedX ? sizeof(double) : 0))
               : 0;
}

template <class F, class R, int MODE>
inline cudaError_t launch_one(const Plan& P, const DeviceState& st, const LaunchConfig& cfg, cudaStream_t stream) {
    QuadParams<F, R, MODE> prm;
    prm.P = P;
    prm.st = st;
    if constexpr (PartWorker<F, R, MODE>::kUniformRows) fill_row_consts(P, &prm.rc);
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(cfg.grid, 1, 1);
    lc.blockDim = dim3(kThreads, 1, 1);
    lc.dynamicSmemBytes = rotate_smem_bytes<F, R, MODE>();
    lc.stream = stream;
    if (lc.dynamicSmemBytes > 48 * 1024) {
        // above the default limit (only the double-precision x-dependent trig integrand): opt in,
        // once per device
        static std::atomic<bool> done[64];   // zero-initialised; set once per device, any thread
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && !done[dev].load(std::memory_order_acquire)) {
            cudaError_t e = cudaFuncSetAttribute(quad_kernel<F, R, MODE>,
                                                 cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 int(lc.dynamicSmemBytes));
            if (e != cudaSuccess) return e;
            done[dev].store(true, std::memory_order_release);
        }
    }
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cfg.cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    lc.attrs = attr;
    lc.numAttrs = 1;
    return cudaLaunchKernelEx(&lc, quad_kernel<F, R, MODE>, prm);
}

}  // namespace bq
