// Per-point / per-thread arithmetic shared verbatim by the sm_100a kernel
// (quad_kernel.cu) and by the CPU backend (cpu_backend.cpp), so both walk the
// same summation tree.  Nothing here touches memory.
#pragma once
#include "bq_dd.cuh"
#include "integrands.cuh"
#include "plan.h"

namespace bq {

// sin and cos of one angle.
struct SinCos {
    double s, c;
};

// sin/cos of (angle of `base` + angle of `rot`): 2 DMUL + 2 DFMA.
BQ_HD SinCos rotate_sc(SinCos base, SinCos rot) {
    return SinCos{fma(base.c, rot.s, base.s * rot.c), fma(-base.s, rot.s, base.c * rot.c)};
}

// ROTATE mode builds every sample's sin/cos from accurately computed anchors by
// exact angle sums, never by a recurrence longer than two rotations:
//   tile base      sincos(angle(x_tile))                 libm/libdevice, once per tile
//   row base       rotate(tile base, sincos(angle(r * kWidth * h)))
//   column const   rotate(sincos(angle(t * h)), sincos(angle(j * kThreads * h)))
//   sample         rotate(row base, column const)        (point_rotate below)
// ROTATE: f at angle(x0) + angle(dx) from the row base (S, C) = sincos(angle(x0))
// and the column constants (cj, sj) = cos/sin(angle(dx)).  DMUL + DFMA per
// needed component; for SinF that is the whole evaluation.
template <class F, class R>
BQ_HD R point_rotate(const F& f, R S, R C, R cj, R sj, double x0, double dx) {
    R s = R(0), c = R(0);
    if (F::kNeedS) s = bq_fma(C, sj, S * cj);
    if (F::kNeedC) c = bq_fma(-S, sj, C * cj);
    R x = F::kNeedX ? R(x0 + dx) : R(0);
    return f.eval_sc(x, s, c);
}

// DIRECT: x = a + (i + off) * h with a single rounding (the index arithmetic
// rowbase + col is exact in fp64 below 2^52), then a plain call of f.
// FAST: 0 = libm / libdevice; 1 = the plan has verified the argument range of the bounded sin/cos
// (kFlagFastTrig); 2 = as 1, and the samples of a row share one multiple of pi/2 (kFlagRowK; only
// whole rows differ, a single point is evaluated as in 1).
template <class F, class R, int FAST = 0>
BQ_HD R point_direct(const F& f, double rowbase, double col, double h, double a) {
    double x = fma(rowbase + col, h, a);
    if constexpr (FAST != 0) return f.template eval_fast<R>(R(x));
    return f.template eval<R>(R(x));
}

// DIRECT: the kCols samples of one row of one thread (columns kThreads points apart).
template <class F, class R, int FAST>
struct RowDirect {
    static BQ_HD void eval(const F& f, double rowbase, const double (&col)[kCols], double h, double a,
                           R (&v)[kCols]) {
#pragma unroll
        for (int j = 0; j < kCols; ++j) v[j] = point_direct<F, R, FAST>(f, rowbase, col[j], h, a);
    }
};
// Trig integrands in fp64 with a verified argument range: bounded reduction per sample, then the
// polynomial.  On the GPU a warp-row whose 8 x 32 samples all fall in quadrants of one parity
// (nearly always: neighbouring samples are a tiny angle apart) runs the unselected sine or cosine
// polynomial — the selected form costs 20 select instructions per sample, more issue slots than
// its 12 fp64 operations; both forms produce the same bits (integrands.cuh).
template <class F, bool ROWK>
struct RowDirectTrig {
    static BQ_HD void eval(const F& f, double rowbase, const double (&col)[kCols], double h, double a,
                           double (&v)[kCols]) {
        constexpr bool COS = F::kNeedC;
        double x[kCols];
        TrigRed red[kCols];
#pragma unroll
        for (int j = 0; j < kCols; ++j) {
            x[j] = fma(rowbase + col[j], h, a);
            if (ROWK && j > 0)
                red[j] = trig_reduce_with<COS>(f.angle_fast(x[j]), red[0]);
            else
                red[j] = trig_reduce<COS>(f.angle_fast(x[j]));
        }
        double val[kCols];
        bool done = false;
#if defined(__CUDA_ARCH__)
        unsigned odd = 0;
        if (ROWK) {
            odd = (red[0].q & 1) ? (1u << kCols) - 1u : 0u;     // one quadrant for the whole row
        } else {
#pragma unroll
            for (int j = 0; j < kCols; ++j) odd |= unsigned(red[j].q & 1) << j;
        }
        if (__all_sync(0xffffffffu, odd == 0u)) {
#pragma unroll
            for (int j = 0; j < kCols; ++j) val[j] = trig_poly_sin(red[j].r);
            done = true;
        } else if (__all_sync(0xffffffffu, odd == (1u << kCols) - 1u)) {
#pragma unroll
            for (int j = 0; j < kCols; ++j) val[j] = trig_poly_cos(red[j].r);
            done = true;
        }
#endif
        if (!done) {
#pragma unroll
            for (int j = 0; j < kCols; ++j) val[j] = trig_poly_sel(red[j].r, (red[j].q & 1) != 0);
        }
#pragma unroll
        for (int j = 0; j < kCols; ++j) {
            const double sc = trig_sign(val[j], red[j].q);
            v[j] = COS ? f.template eval_sc<double>(x[j], 0.0, sc) : f.template eval_sc<double>(x[j], sc, 0.0);
        }
    }
};
template <>
struct RowDirect<SinF, double, 1> : RowDirectTrig<SinF, false> {};
template <>
struct RowDirect<AccF, double, 1> : RowDirectTrig<AccF, false> {};
template <>
struct RowDirect<VelF, double, 1> : RowDirectTrig<VelF, false> {};
template <>
struct RowDirect<DisF, double, 1> : RowDirectTrig<DisF, false> {};
template <>
struct RowDirect<SinF, double, 2> : RowDirectTrig<SinF, true> {};
template <>
struct RowDirect<AccF, double, 2> : RowDirectTrig<AccF, true> {};
template <>
struct RowDirect<VelF, double, 2> : RowDirectTrig<VelF, true> {};
template <>
struct RowDirect<DisF, double, 2> : RowDirectTrig<DisF, true> {};

// The table integrand: the row is a run of equally spaced samples, so when its two ends
// share a table cell (nearly always) each sample is one fma (ProfileF::eval_run).
template <class R, int FAST>
struct RowDirect<ProfileF, R, FAST> {
    static BQ_HD void eval(const ProfileF& f, double rowbase, const double (&col)[kCols], double h, double a,
                           R (&v)[kCols]) {
        const double x_first = fma(rowbase + col[0], h, a);
        const double x_last = fma(rowbase + col[kCols - 1], h, a);
        if (f.template eval_run<kCols>(x_first, x_last, v)) return;
#pragma unroll
        for (int j = 0; j < kCols; ++j) v[j] = point_direct<ProfileF, R, 0>(f, rowbase, col[j], h, a);
    }
};

// Fused accumulation (integrands with kFusedSum, ROTATE, both precisions).  For f = alpha*s + beta*c + gamma
// with s = S*cj + C*sj, c = C*cj - S*sj the sample is S*P + C*Q + gamma with the per-column
// constants P = alpha*cj - beta*sj, Q = alpha*sj + beta*cj, so the loop is two chained FMAs,
// acc = fma(C, Q, fma(S, P, acc)), whatever the linear combination; gamma enters once, as
// gamma * n, in finish_result.  (alpha, beta) = (1, 0) leaves cj, sj unchanged bit for bit.
template <class F>
BQ_HD void fused_consts(const F& f, SinCos col, double& P, double& Q) {
    const double al = f.fuse_alpha(), be = f.fuse_beta();
    P = al * col.c - be * col.s;
    Q = al * col.s + be * col.c;
}

// fp32 only: the fused loop runs on the integrand divided by its amplitude, so that the column constant
// of the large stream is cos(column offset) — within 8e-6 of 1 on a fine grid, where fused_hi() makes it
// exactly 1.0f and the products with the gridded row values exact (RowDiffusion below).
//   beta == 0:  f / alpha = sin(row + col) = S * cos(col) + C * sin(col)      streams (S, C)
//   alpha == 0: f / beta  = cos(row + col) = C * cos(col) + S * (-sin(col))   streams (C, S): swapped
// and the flush multiplies by the amplitude in fp64.  A genuine two-term combination is left as it is.
struct FusedNorm {
    double amp;
    bool swap;
};
template <class F, class R>
BQ_HD FusedNorm fused_norm(const F& f) {
    if (sizeof(R) == 8) return FusedNorm{1.0, false};
    const double al = f.fuse_alpha(), be = f.fuse_beta();
    if (be == 0.0 && al != 0.0) return FusedNorm{al, false};
    if (al == 0.0 && be != 0.0) return FusedNorm{be, true};
    return FusedNorm{1.0, false};
}
template <class F>
BQ_HD void fused_consts(const F& f, const FusedNorm& nm, SinCos col, double& P, double& Q) {
    if (nm.amp == 1.0 && !nm.swap) {
        fused_consts(f, col, P, Q);
    } else {
        P = col.c;
        Q = nm.swap ? -col.s : col.s;
    }
}

// Per-thread, per-tile accumulator.
//   fp64: kCols independent naive sums (<= kMaxRows terms each), combined
//         pairwise and pushed into the thread's dd total at the end of a tile.
//   fp32: the kCols values of a row are summed pairwise, the row sums are
//         Kahan-accumulated in fp32, and the compensated value is promoted to
//         fp64 at the end of a tile.
// Fixed pairwise (balanced-tree) sum of the kCols values of one row.
template <class R>
BQ_HD R pairwise_cols(const R (&v)[kCols]) {
    static_assert(kCols == 8, "pairwise_cols is written for 8 columns");
    return ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
}

template <class R>
struct TileAcc;

template <>
struct TileAcc<double> {
    double a[kCols];
    BQ_HD void clear() {
#pragma unroll
        for (int j = 0; j < kCols; ++j) a[j] = 0.0;
    }
    BQ_HD void add_row(const double (&v)[kCols]) {
#pragma unroll
        for (int j = 0; j < kCols; ++j) a[j] += v[j];
    }
    BQ_HD void add_one(int j, double v) { a[j] += v; }
    // Fused accumulation of the sample S*cj + C*sj (integrands with kFusedSum): two chained FMAs.
    BQ_HD void add_rot(int j, double S, double C, double cj, double sj) { a[j] = fma(C, sj, fma(S, cj, a[j])); }
    // one whole row (kCols samples of this thread)
    BQ_HD void add_rot_row(double S, double C, const double (&cj)[kCols], const double (&sj)[kCols]) {
#pragma unroll
        for (int j = 0; j < kCols; ++j) add_rot(j, S, C, cj[j], sj[j]);
    }
    BQ_HD void promote(bool) {}
    BQ_HD double flush() const { return pairwise_cols<double>(a); }
    BQ_HD double flush_fused(bool) { return pairwise_cols<double>(a); }
};

template <>
struct TileAcc<float> {
    // sample-by-sample integrands: Kahan in fp32 over the pairwise row sums
    float sum, comp;
    // fused integrands: fp32 FMA runs of at most kPromoteRows rows per column, promoted to fp64
    // (the fp64 adds ride on the otherwise idle fp64 pipe).  The two products of a sample go into
    // two accumulators per column: for a fine grid the C*Q terms (Q = sin(column * h)) are below half
    // an ulp of the running S*P sum and a single accumulator drops them one by one — a bias, not a
    // random rounding; kept apart they add up among themselves and meet the other sum once per run.
    float a[kCols], c[kCols];
    double d[kCols];
    BQ_HD void clear() {
        sum = 0.f, comp = 0.f;
#pragma unroll
        for (int j = 0; j < kCols; ++j) a[j] = 0.f, c[j] = 0.f, d[j] = 0.0;
    }
    BQ_HD void add_row(const float (&v)[kCols]) {
        kahan_add(sum, comp, pairwise_cols<float>(v));
    }
    BQ_HD void add_one(int, float v) { kahan_add(sum, comp, v); }
    BQ_HD void add_rot(int j, float S, float C, float P, float Q) {
        a[j] = fmaf(S, P, a[j]);
        c[j] = fmaf(C, Q, c[j]);
    }
    // One whole row.  On the GPU the columns go two at a time through the packed fp32 FMA
    // (fma.rn.f32x2 -> SASS FFMA2, Blackwell).  Each half is an IEEE fma, so the bits equal the
    // scalar form the CPU tree uses.
    BQ_HD void add_rot_row(float S, float C, const float (&P)[kCols], const float (&Q)[kCols]) {
#if defined(__CUDA_ARCH__)
        static_assert(kCols % 2 == 0, "columns are paired");
        unsigned long long s2, c2;
        asm("mov.b64 %0, {%1, %1};" : "=l"(s2) : "f"(S));
        asm("mov.b64 %0, {%1, %1};" : "=l"(c2) : "f"(C));
#pragma unroll
        for (int j = 0; j < kCols; j += 2) {
            unsigned long long a2, b2, p2, q2;
            asm("mov.b64 %0, {%1, %2};" : "=l"(a2) : "f"(a[j]), "f"(a[j + 1]));
            asm("mov.b64 %0, {%1, %2};" : "=l"(b2) : "f"(c[j]), "f"(c[j + 1]));
            asm("mov.b64 %0, {%1, %2};" : "=l"(p2) : "f"(P[j]), "f"(P[j + 1]));
            asm("mov.b64 %0, {%1, %2};" : "=l"(q2) : "f"(Q[j]), "f"(Q[j + 1]));
            asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(a2) : "l"(s2), "l"(p2));
            asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(b2) : "l"(c2), "l"(q2));
            asm("mov.b64 {%0, %1}, %2;" : "=f"(a[j]), "=f"(a[j + 1]) : "l"(a2));
            asm("mov.b64 {%0, %1}, %2;" : "=f"(c[j]), "=f"(c[j + 1]) : "l"(b2));
        }
#else
#pragma unroll
        for (int j = 0; j < kCols; ++j) add_rot(j, S, C, P[j], Q[j]);
#endif
    }
    // End of a run of kPromoteRows rows: the S*P sum of a column joins the column's fp64 sum.  The C*Q
    // sum follows it at the end of the flush group on a fine grid (kFlagFineGrid: it stays tiny, and
    // an fp32 add of the two would round it to the ulp of the large one — a bias of 3e-10 on [0, pi/2]
    // at 2^30); on a coarse grid the two are of one size and meet in an fp32 add at every run.
    // (A packed add.f32x2 for that add measured slower than eight scalar adds.)
    BQ_HD void promote(bool fine) {
        if (fine) {
#pragma unroll
            for (int j = 0; j < kCols; ++j) d[j] += double(a[j]), a[j] = 0.f;
        } else {
#pragma unroll
            for (int j = 0; j < kCols; ++j) d[j] += double(a[j] + c[j]), a[j] = 0.f, c[j] = 0.f;
        }
    }
    BQ_HD double flush() const { return double(sum) - double(comp); }
    BQ_HD double flush_fused(bool fine) {
        promote(fine);
        if (fine) {
#pragma unroll
            for (int j = 0; j < kCols; ++j) d[j] += double(c[j]), c[j] = 0.f;
        }
        return pairwise_cols<double>(d);
    }
};

// Row values of the fused fp32 path, rounded with error diffusion.
//
// A column's fp32 accumulator adds 16 or 32 row values S_r (24 significant bits, |S_r| <= 1) one after
// the other; past |sum| = 2 every add rounds.  On a fine grid all 2048 columns of a row carry the
// same S_r * P_j with P_j = cos(column * h) = 1.0f, so those roundings are the same in every column
// and change slowly from run to run: they do not average out (measured: 1.9e-10 at n = 2^32 with
// 16-row runs against 2.8e-11 with a promotion after every row, 3.8e-13 with the diffusion).  So the row values are put on a
// grid of 2^-18 instead — S~ = rint((S + e) * 2^18) / 2^18, e = what the earlier rows of this lane's
// row stream still owe — which makes every partial sum of a run (< 64) exact in fp32 whenever
// P_j = 1.0f, and never worse than before otherwise.  The grid error does not accumulate: each
// lane carries its e from tile to tile, and what is owed at the end of a part is added to the sums
// once, in fp64 (diffusion_owed).  Full rows only (every column uses them): all but the grid's ragged last row.
template <class R, bool FUSED>
struct RowDiffusion {
    static constexpr bool kOn = false;
    double es = 0.0, ec = 0.0, ss = 0.0, sc = 0.0;
    BQ_HD void round(const SinCos& rb, bool, R& S, R& C) { S = R(rb.s), C = R(rb.c); }
};
template <>
struct RowDiffusion<float, true> {
    static constexpr bool kOn = true;
    double es = 0.0, ec = 0.0;     // what this lane's row stream owes the sums
    double ss = 0.0, sc = 0.0;     // sums of the stream's fp32 row values (for the column constants' fp32 rounding)
    static BQ_HD float grid(double t, double& owed) {
        constexpr double kMagic = 0x1.8p+34;            // ulp = 2^-18: (t + magic) - magic = t rounded to the grid
        const double q = (t + kMagic) - kMagic;         // |q| <= 1 + 2^-18, 20 significant bits: float(q) is exact
        owed = t - q;
        return float(q);
    }
    // full_row: every column uses the row (all rows but the grid's ragged last one).
    BQ_HD void round(const SinCos& rb, bool full_row, float& S, float& C) {
        if (full_row) {
            S = grid(rb.s + es, es);
            C = grid(rb.c + ec, ec);
            ss += double(S), sc += double(C);
        } else {
            S = float(rb.s), C = float(rb.c);
        }
    }
};
// The fp32 column constant that multiplies the S stream.  On a fine grid P = cos(column * h * w) sits
// within 8e-6 of 1: it is taken as exactly 1.0f there (products with the gridded row values are then
// exact) and the whole difference goes into the fp64 closing term below, instead of the nearest fp32
// value 1 - k * 2^-24 whose products round, in one direction, in every row.
template <class R>
BQ_HD R fused_hi(double Pc, bool fine) {
    if (sizeof(R) == 4 && fine && fabs(Pc - 1.0) <= 0x1p-16) return R(1);
    return R(Pc);
}

// The part's closing term of row stream r (added by lane r of the CTA's first warp), in fp64:
//   (what the stream owes in S) * PS + (what it owes in C) * QS                   the grid residues
// + (its sum of S) * PLO + (its sum of C) * QLO     the fp32 rounding of the column constants, which is the
//                                                   same in every row and so a bias (7.5e-9 at n = 2^24 without)
// PS, QS, PLO, QLO: sums over the CTA's 2048 columns of float(P_j), float(Q_j), P_j - float(P_j),
// Q_j - float(Q_j), formed once per launch.
BQ_HD double diffusion_owed(double es, double ec, double ss, double sc, double PS, double QS, double PLO,
                            double QLO) {
    return fma(es, PS, ec * QS) + fma(ss, PLO, sc * QLO);
}

#ifndef BQ_PROMOTE_ROWS
#define BQ_PROMOTE_ROWS 32
#endif
constexpr int kPromoteRows = BQ_PROMOTE_ROWS;   // fused fp32: rows summed in fp32 before the promotion to fp64

// Geometry of one tile.
struct TileGeom {
    uint64_t start;      // first point index
    uint32_t full_rows;  // rows with kWidth points
    uint32_t rem;        // points in the ragged last row (0 = none)
};

BQ_HDI TileGeom tile_geom(const Plan& P, uint64_t tile) {
    TileGeom g;
    g.start = tile * P.tile_pts;
    uint64_t left = P.n - g.start;
    uint64_t npts = left < P.tile_pts ? left : P.tile_pts;
    g.full_rows = uint32_t(npts / kWidth);
    g.rem = uint32_t(npts % kWidth);
    return g;
}

// First-index-as-double of row r of a tile (exact), including the rule offset.
BQ_HDI double row_base_index(const Plan& P, const TileGeom& g, uint32_t r) {
    return double(g.start + uint64_t(r) * kWidth) + P.off;
}

template <class F>
struct FunctorOf;

BQ_HDI SinF make_sin(const Plan&) { return SinF{}; }
BQ_HDI AccF make_acc(const Plan& P) { return AccF{P.p[0], P.p[1], 1.0 / P.p[0]}; }
BQ_HDI VelF make_vel(const Plan& P) { return VelF{P.p[0], P.p[2], 1.0 / P.p[0]}; }
BQ_HDI DisF make_dis(const Plan& P) { return DisF{P.p[0], P.p[2], 1.0 / P.p[0]}; }
BQ_HDI ProfileF make_profile(const Plan& P, const double* table) {
    const double inv_dx = 1.0 / P.p[0];
    // a row's samples are kThreads points apart (quad kernel); the scan overrides step_du
    return ProfileF{table, int(P.table_len), inv_dx, double(kThreads) * P.h * inv_dx};
}

template <class F>
BQ_HDI F make_functor(const Plan& P, const double* table);
template <>
BQ_HDI SinF make_functor<SinF>(const Plan& P, const double*) { return make_sin(P); }
template <>
BQ_HDI AccF make_functor<AccF>(const Plan& P, const double*) { return make_acc(P); }
template <>
BQ_HDI VelF make_functor<VelF>(const Plan& P, const double*) { return make_vel(P); }
template <>
BQ_HDI DisF make_functor<DisF>(const Plan& P, const double*) { return make_dis(P); }
template <>
BQ_HDI ProfileF make_functor<ProfileF>(const Plan& P, const double* t) { return make_profile(P, t); }

// Last stage, shared by kernel / CPU / bq_fold_segments: fold the 8 segment
// pairs in index order, scale by h, add the trapezoid end correction
// (trap = left + h/2 (f(b) - f(a)); the reference has no trapezoid, only the
// commented-out include at 4main.c:38).
// sum over the grid of x_i = a + (i + off) * h, i < n:  n * a + h * (n (n - 1) / 2 + n * off), as a pair.
// n < 2^53 and off is 0, 1/2 or 1, so every factor below is exact and the two products are exact pairs.
BQ_HDI dd grid_sum_x(const Plan& P) {
    const double n = double(P.n);
    const bool odd = (P.n & 1ull) != 0ull;                  // one of n, n - 1 is even: halve that one
    const double u = odd ? n : n * 0.5;
    const double v = odd ? double(P.n - 1) * 0.5 : double(P.n - 1);
    dd tri = dd_mul_d(dd{u, 0.0}, v);                       // n (n - 1) / 2
    tri = dd_add_d(tri, n * P.off);
    return dd_add(dd_mul_d(tri, P.h), dd_mul_d(dd{n, 0.0}, P.a));
}

template <class F>
BQ_HDI void finish_result(const Plan& P, const F& f, const double* seg_hi, const double* seg_lo,
                          bq_result* out) {
    dd acc{0.0, 0.0};
    for (int s = 0; s < kSegments; ++s) acc = dd_add(acc, dd{seg_hi[s], seg_lo[s]});
    if (F::kFusedSum && P.eval == BQ_EVAL_ROTATE && f.fuse_gamma() != 0.0)
        acc = dd_add(acc, dd_mul_d(dd{double(P.n), 0.0}, f.fuse_gamma()));   // the constant term of every sample
    if constexpr (F::kFusedSum) {
        if (P.eval == BQ_EVAL_ROTATE && f.fuse_delta() != 0.0)
            acc = dd_add(acc, dd_mul_d(grid_sum_x(P), f.fuse_delta()));      // the linear term of every sample
        if (P.eval == BQ_EVAL_ROTATE && f.fuse_scale() != 1.0) acc = dd_mul_d(acc, f.fuse_scale());
    }
    dd val = dd_mul_d(acc, P.h);
    if (P.rule == BQ_RULE_TRAP) {
        const bool fast = (P.flags & kFlagFastTrig) != 0;   // same sin/cos as the samples: host == device bits
        double fb = fast ? f.template eval_fast<double>(P.b) : f.template eval<double>(P.b);
        double fa = fast ? f.template eval_fast<double>(P.a) : f.template eval<double>(P.a);
        val = dd_add_d(val, 0.5 * P.h * (fb - fa));
    }
    out->raw_sum_hi = acc.hi;
    out->raw_sum_lo = acc.lo;
    out->value = val.hi;
    out->value_lo = val.lo;
    for (int s = 0; s < kSegments; ++s) {
        out->seg_hi[s] = seg_hi[s];
        out->seg_lo[s] = seg_lo[s];
    }
    out->n = P.n;
    out->eval_used = P.eval;
    out->world = P.world;
    out->rank = P.rank;
    out->launches = 1;
}

}  // namespace bq
