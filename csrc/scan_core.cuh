// Arithmetic of the fused scan shared by scan_kernel.cu and cpu_scan.cpp.
//
// A run of samples is summarised by (p, q, cnt):
//   p   = sum of the values v_k             ("default_sum" increment, 4main.c:105)
//   q   = sum of the local running sums     ("default_sum_of_sums" increment, 4main.c:185)
//   cnt = number of samples
// and two adjacent runs A (earlier), B (later) combine as
//   p = pA + pB,  q = qA + qB + cntB * pA,  cnt = cntA + cntB
// which is what the reference's root-side carry fix-up (4main.c:147-154,
// 211-218) computes one slice at a time.
#pragma once
#include "bq_dd.cuh"
#include "integrands.cuh"
#include "scan_plan.h"

namespace bq {

struct Run {          // dd precision: used across tiles / segments
    dd p, q;
    double cnt;       // exact (integers below 2^53)
};

BQ_HD Run run_identity() { return Run{dd{0.0, 0.0}, dd{0.0, 0.0}, 0.0}; }

BQ_HD Run run_combine(const Run& A, const Run& B) {
    Run r;
    r.p = dd_add(A.p, B.p);
    r.q = dd_add(dd_add(A.q, B.q), dd_mul_d(A.p, B.cnt));
    r.cnt = A.cnt + B.cnt;
    return r;
}

BQ_HD Run run_from_tile(double s, double w, double cnt) {
    return Run{dd{s, 0.0}, dd{w, 0.0}, cnt};
}

struct Run2 {         // plain fp64: used inside one tile
    double p, q;
};

// A earlier, B later covering cntB samples.
BQ_HD Run2 run2_combine(const Run2& A, const Run2& B, double cntB) {
    return Run2{A.p + B.p, fma(cntB, A.p, A.q + B.q)};
}

BQ_HDI uint32_t scan_tile_count(const ScanPlan& P, uint64_t tile) {
    const uint64_t start = tile * uint64_t(kScanTile);
    const uint64_t left = P.n - start;
    return left < uint64_t(kScanTile) ? uint32_t(left) : uint32_t(kScanTile);
}

// Samples covered by tiles [a, b): every tile is full except the last one of the grid.
BQ_HDI double scan_samples_in_tiles(const ScanPlan& P, uint64_t a, uint64_t b) {
    if (b <= a) return 0.0;
    const uint64_t first = a * uint64_t(kScanTile);
    uint64_t last = b * uint64_t(kScanTile);
    if (last > P.n) last = P.n;
    return double(last - first);
}

BQ_HDI int scan_segment_of(const ScanPlan& P, uint64_t tile) {
    int s = 0;
    for (int k = 1; k < kSegments; ++k) s += (tile >= P.seg_tile[k]) ? 1 : 0;
    return s;
}

// The integrand functor of a scan: built from the quad plan fields it reads (parameters, table
// length); for the table integrand the run step is the grid spacing, one sample apart.
template <class F>
BQ_HDI F make_scan_functor(const ScanPlan& P, const double* table) {
    Plan fp{};
    for (int i = 0; i < 4; ++i) fp.p[i] = P.p[i];
    fp.table_len = P.table_len;
    fp.h = P.dx / double(kThreads);   // make_profile: step_du = kThreads * h / spacing = dx / spacing (exact: a power of two)
    return make_functor<F>(fp, table);
}

template <class F>
struct ScanRun {   // the run of kScanItems consecutive samples starting at index `base`; false = not handled
    static BQ_HD bool eval(const F&, const ScanPlan&, double, double (&)[kScanItems]) { return false; }
};
template <>
struct ScanRun<ProfileF> {
    static BQ_HD bool eval(const ProfileF& f, const ScanPlan& P, double base, double (&v)[kScanItems]) {
        return f.template eval_run<kScanItems>(fma(base, P.dx, P.x0), fma(base + double(kScanItems - 1), P.dx, P.x0), v);
    }
};

// Thread-level work on its kScanItems consecutive samples of a tile.
template <class F>
BQ_HD void scan_thread_values(const F& f, const ScanPlan& P, uint64_t first_idx, uint32_t k0,
                              uint32_t m, double (&v)[kScanItems]) {
    // One integer -> double conversion per run; base + j is exact below 2^53, so
    // this is double(first_idx + j) bit for bit.
    const double base = double(first_idx);
    if (k0 + uint32_t(kScanItems) <= m) {
        // table integrand: cell and slope once per run, one fma per sample (ProfileF::eval_run)
        if (ScanRun<F>::eval(f, P, base, v)) return;
        // The whole run lies inside the grid (every run but those of the last tile):
        // straight-line code, so the kScanItems evaluations overlap.  A per-sample
        // guard here turns into one branch per sample and serialises them.
        if (F::kTrig && (P.flags & kScanFastTrig)) {     // uniform: a kernel parameter
#pragma unroll
            for (int j = 0; j < kScanItems; ++j) v[j] = f.template eval_fast<double>(fma(base + double(j), P.dx, P.x0));
            return;
        }
#pragma unroll
        for (int j = 0; j < kScanItems; ++j) v[j] = f.template eval<double>(fma(base + double(j), P.dx, P.x0));
    } else {
        const bool fast = F::kTrig && (P.flags & kScanFastTrig);
#pragma unroll
        for (int j = 0; j < kScanItems; ++j) {
            const double x = fma(base + double(j), P.dx, P.x0);
            v[j] = (k0 + uint32_t(j) < m) ? (fast ? f.template eval_fast<double>(x) : f.template eval<double>(x)) : 0.0;
        }
    }
}

// A part: (sum v, sum v*(m-k)) of this thread's samples.  For a whole run the weights are taken out
// of the loop, sum_j v_j (m - k0 - j) = (m - k0) * sum_j v_j - sum_j j * v_j: one integer->double
// conversion per run instead of one per sample (the conversions run on a quarter-rate unit and were
// the bound of this part: 8 of them cost more than all its fp64 arithmetic).
BQ_HD void scan_thread_aggregate(const double (&v)[kScanItems], uint32_t k0, uint32_t m, double& s,
                                 double& w) {
    s = 0.0;
    if (k0 + uint32_t(kScanItems) <= m) {
        double t = 0.0;
#pragma unroll
        for (int j = 0; j < kScanItems; ++j) {
            s += v[j];
            if (j) t = fma(double(j), v[j], t);
        }
        w = fma(double(m - k0), s, -t);
        return;
    }
    w = 0.0;
#pragma unroll
    for (int j = 0; j < kScanItems; ++j) {
        const uint32_t k = k0 + uint32_t(j);
        const double wt = k < m ? double(m - k) : 0.0;
        s += v[j];
        w = fma(v[j], wt, w);
    }
}

// Phase B: local inclusive scans lp_j = sum_{i<=j} v_i, lq_j = sum_{i<=j} lp_i.
BQ_HD void scan_thread_local(const double (&v)[kScanItems], double (&lp)[kScanItems],
                             double (&lq)[kScanItems]) {
    double p = 0.0, q = 0.0;
#pragma unroll
    for (int j = 0; j < kScanItems; ++j) {
        p += v[j];
        q += p;
        lp[j] = p;
        lq[j] = q;
    }
}

// Tile exclusive prefix from the segment base and the in-segment prefix.
BQ_HD Run2 scan_tile_exclusive(const Run& base, const Run& pre) {
    const Run r = run_combine(base, pre);
    return Run2{dd_to_double(r.p), dd_to_double(r.q)};
}

}  // namespace bq
