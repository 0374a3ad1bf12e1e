// CPU backend of cumulative integration: the tile / segment tree of
// scan_kernel.cu on host threads (capability of 4main.c:76-225).
#include <math.h>
#include <string.h>

#include <atomic>
#include <thread>
#include <vector>

#include "host_common.h"
#include "quad_core.cuh"
#include "scan_core.cuh"

namespace bq {

int make_scan_plan(const bq_scan_job& job, int rank, int world, int table_len, ScanPlan* P) {
    memset(P, 0, sizeof(*P));
    if (job.n < 1 || job.n > kScanMaxN) {
        set_error("cumulative: n=%llu out of range [1, 2^34]", (unsigned long long)job.n);
        return BQ_ERR_INVALID;
    }
    if (world < 1 || world > BQ_MAX_WORLD || (kSegments % world) != 0 || rank < 0 || rank >= world) {
        set_error("world=%d rank=%d: world must be 1, 2, 4 or 8", world, rank);
        return BQ_ERR_INVALID;
    }
    if (job.integrand == BQ_F_CUSTOM) {
        set_error("cumulative: custom integrands are supported by integrate() only");
        return BQ_ERR_UNSUPPORTED;
    }
    if (job.integrand < 0 || job.integrand >= BQ_F_COUNT || !isfinite(job.x0) || !isfinite(job.dx)) {
        set_error("cumulative: bad integrand / grid");
        return BQ_ERR_INVALID;
    }
    P->n = job.n;
    P->x0 = job.x0;
    P->dx = job.dx;
    P->integrand = job.integrand;
    for (int i = 0; i < 4; ++i) P->p[i] = job.p[i];
    if (job.integrand == BQ_F_ACC || job.integrand == BQ_F_VEL || job.integrand == BQ_F_DIS) {
        if (P->p[0] == 0.0) P->p[0] = kTscale;
        if (P->p[1] == 0.0) P->p[1] = kAscale;
        if (P->p[2] == 0.0) P->p[2] = kVscale;
    }
    if (job.integrand == BQ_F_PROFILE) {
        if (P->p[0] == 0.0) P->p[0] = 1.0;
        if (table_len < 2 || table_len > kMaxTable) {
            set_error("PROFILE table length %d out of range", table_len);
            return BQ_ERR_INVALID;
        }
        P->table_len = uint32_t(table_len);
        const double xlast = job.x0 + double(job.n - 1) * job.dx;
        const double lo = job.x0 < xlast ? job.x0 : xlast, hi = job.x0 < xlast ? xlast : job.x0;
        const double xmax = double(table_len - 1) * P->p[0];
        if (lo < 0.0 || hi > xmax * (1.0 + 1e-15)) {
            set_error("PROFILE integrand is defined on [0, %g], grid spans [%g, %g]", xmax, lo, hi);
            return BQ_ERR_RANGE;
        }
    }
    P->rank = rank;
    P->world = world;
    P->flags = (job.replicate ? kScanReplicate : 0u) | (job.want_samples ? kScanSamples : 0u);
    if (job.integrand == BQ_F_SIN || job.integrand == BQ_F_ACC || job.integrand == BQ_F_VEL || job.integrand == BQ_F_DIS) {
        // trig integrands: the bounded-argument sin/cos applies when the whole grid stays inside its range
        const double w = job.integrand == BQ_F_SIN ? 1.0 : 1.0 / fabs(P->p[0]);
        const double xl = job.x0 + double(job.n) * job.dx;
        const double amax = fabs(job.x0) > fabs(xl) ? fabs(job.x0) : fabs(xl);
        if (w * amax <= kFastTrigMax) P->flags |= kScanFastTrig;
    }
    P->timeout_ms = 2000;
    P->n_tiles = (job.n + kScanTile - 1) / kScanTile;
    for (int s = 0; s <= kSegments; ++s) P->seg_tile[s] = (uint64_t(s) * P->n_tiles) / kSegments;
    P->seg_group0[0] = P->seg_super0[0] = 0;
    for (int s = 0; s < kSegments; ++s) {
        const uint64_t T = P->seg_tile[s + 1] - P->seg_tile[s];
        const uint64_t G = (T + kScanGroup - 1) / kScanGroup;
        P->seg_group0[s + 1] = P->seg_group0[s] + G;
        P->seg_super0[s + 1] = P->seg_super0[s] + (G + kScanSuper - 1) / kScanSuper;
    }
    const uint32_t segs = kSegments / world;
    P->seg_lo = rank * segs;
    P->seg_hi = P->seg_lo + segs;
    P->tile_lo = P->seg_tile[P->seg_lo];
    P->tile_hi = P->seg_tile[P->seg_hi];
    return BQ_OK;
}

namespace {

// Kogge-Stone inclusive scan over 32 "lanes", as the kernel's shuffle loops do.
template <class T, class Op>
void lanes_inclusive(T* x, Op op) {
    for (int d = 1; d < 32; d <<= 1) {
        T tmp[32];
        for (int l = 0; l < 32; ++l) tmp[l] = l >= d ? op(x[l - d], x[l], d) : x[l];
        for (int l = 0; l < 32; ++l) x[l] = tmp[l];
    }
}

double lanes_sum(const double* v, int lanes) {
    double x[32], t[32];
    for (int l = 0; l < lanes; ++l) x[l] = v[l];
    for (int off = lanes / 2; off >= 1; off >>= 1) {
        for (int l = 0; l < lanes; ++l) t[l] = x[l] + x[l ^ off];
        for (int l = 0; l < lanes; ++l) x[l] = t[l];
    }
    return x[0];
}

template <class F>
struct CpuScan {
    const ScanPlan& P;
    F f;
    std::vector<double> agg, prefix;     // global tile indexing
    Run seg_tot[kSegments];
    Run seg_base[kSegments];

    CpuScan(const ScanPlan& P_, const double* table) : P(P_) {
        f = make_scan_functor<F>(P, table);
        agg.assign(2 * P.n_tiles, 0.0);
        prefix.assign(4 * P.n_tiles, 0.0);
    }

    void tile_aggregate(uint64_t tile) {
        const uint32_t m = scan_tile_count(P, tile);
        // thread (pass w, lane l) -> per-lane fold over the 8 passes in the order
        // ((0+4)+(2+6)) + ((1+5)+(3+7)) -> xor butterfly over the 32 lanes
        constexpr int kPasses = kScanThreads / 32;
        static_assert(kPasses == 8, "the pass fold below is written for 8 passes");
        double ps[kPasses][32], pw[kPasses][32];
        for (int w = 0; w < kPasses; ++w) {
            for (int l = 0; l < 32; ++l) {
                const uint32_t k0 = uint32_t(w * 32 + l) * kScanItems;
                double v[kScanItems];
                scan_thread_values(f, P, tile * uint64_t(kScanTile) + k0, k0, m, v);
                scan_thread_aggregate(v, k0, m, ps[w][l], pw[w][l]);
            }
        }
        double ls[32], lw[32];
        for (int l = 0; l < 32; ++l) {
            ls[l] = ((ps[0][l] + ps[4][l]) + (ps[2][l] + ps[6][l])) + ((ps[1][l] + ps[5][l]) + (ps[3][l] + ps[7][l]));
            lw[l] = ((pw[0][l] + pw[4][l]) + (pw[2][l] + pw[6][l])) + ((pw[1][l] + pw[5][l]) + (pw[3][l] + pw[7][l]));
        }
        agg[2 * tile] = lanes_sum(ls, 32);
        agg[2 * tile + 1] = lanes_sum(lw, 32);
    }

    Run tile_run(uint64_t t) const {
        return run_from_tile(agg[2 * t], agg[2 * t + 1], double(scan_tile_count(P, t)));
    }

    // Samples covered by tiles [a, b) (only the last tile of the grid is short).
    double span(uint64_t a, uint64_t b) const { return scan_samples_in_tiles(P, a, b); }

    // The fixed-shape prefix tree of scan_plan.h, segment by segment: what every tile of the kernel
    // assembles for itself from the aggregates of earlier tiles.
    void segment_prefixes(int seg) {
        const uint64_t t0 = P.seg_tile[seg], t1 = P.seg_tile[seg + 1];
        const uint64_t T = t1 - t0;
        const uint64_t G = (T + kScanGroup - 1) / kScanGroup;
        const uint64_t SG = (G + kScanSuper - 1) / kScanSuper;
        auto combine = [](const Run& a, const Run& b, int) { return run_combine(a, b); };
        Run sgp = run_identity();                      // SGP(sg-1)
        for (uint64_t sg = 0; sg < SG; ++sg) {
            Run gts[kScanSuper];
            std::vector<Run> aexcl(size_t(kScanSuper) * kScanGroup, run_identity());
            int gi_last = -1;
            for (int gi = 0; gi < kScanSuper; ++gi) {
                gts[gi] = run_identity();
                const uint64_t g = sg * kScanSuper + gi;
                if (g >= G) continue;
                gi_last = gi;
                Run x[32];
                int k_last = -1;
                for (int l = 0; l < 32; ++l) {
                    const uint64_t j = g * kScanGroup + l;
                    x[l] = j < T ? tile_run(t0 + j) : run_identity();
                    if (j < T) k_last = l;
                }
                lanes_inclusive(x, combine);
                for (int l = 0; l < 32; ++l) aexcl[size_t(gi) * kScanGroup + l] = l ? x[l - 1] : run_identity();
                gts[gi] = x[k_last];                   // GT(g)
            }
            Run gx[32];
            for (int gi = 0; gi < 32; ++gi) gx[gi] = gts[gi];
            lanes_inclusive(gx, combine);              // GTscan(sg)
            for (int gi = 0; gi <= gi_last; ++gi) {
                const Run gexcl = gi ? gx[gi - 1] : run_identity();
                const Run head = run_combine(sgp, gexcl);
                for (int l = 0; l < 32; ++l) {
                    const uint64_t j = (sg * kScanSuper + gi) * kScanGroup + l;
                    if (j >= T) break;
                    const Run pre = run_combine(head, aexcl[size_t(gi) * kScanGroup + l]);
                    double* o = &prefix[4 * (t0 + j)];
                    o[0] = pre.p.hi, o[1] = pre.p.lo, o[2] = pre.q.hi, o[3] = pre.q.lo;
                }
            }
            sgp = run_combine(sgp, gx[gi_last]);       // SGP(sg)
        }
        seg_tot[seg] = sgp;
    }

    void bases() {
        Run base = run_identity();
        for (int s = 0; s < kSegments; ++s) {
            seg_base[s] = base;
            Run t = seg_tot[s];
            const uint64_t first = P.seg_tile[s] * uint64_t(kScanTile);
            uint64_t last = P.seg_tile[s + 1] * uint64_t(kScanTile);
            if (last > P.n) last = P.n;
            t.cnt = double(last - first);
            base = run_combine(base, t);
        }
    }

    void tile_store(uint64_t tile, double* Pout, double* Qout, double* Sout) const {
        const uint32_t m = scan_tile_count(P, tile);
        const int seg = scan_segment_of(P, tile);
        const double* p = &prefix[4 * tile];
        Run pre{dd{p[0], p[1]}, dd{p[2], p[3]}, double((tile - P.seg_tile[seg]) * uint64_t(kScanTile))};
        const Run2 tile_excl = scan_tile_exclusive(seg_base[seg], pre);
        static thread_local double LP[kScanThreads][kScanItems], LQ[kScanThreads][kScanItems],
            V[kScanThreads][kScanItems];
        Run2 inc[kScanThreads], warp_tot[kScanThreads / 32];
        for (int w = 0; w < kScanThreads / 32; ++w) {
            Run2 x[32];
            for (int l = 0; l < 32; ++l) {
                const int t = w * 32 + l;
                const uint32_t k0 = uint32_t(t) * kScanItems;
                scan_thread_values(f, P, tile * uint64_t(kScanTile) + k0, k0, m, V[t]);
                scan_thread_local(V[t], LP[t], LQ[t]);
                x[l] = Run2{LP[t][kScanItems - 1], LQ[t][kScanItems - 1]};
            }
            lanes_inclusive(x, [](const Run2& a, const Run2& b, int d) {
                return run2_combine(a, b, double(kScanItems * d));
            });
            for (int l = 0; l < 32; ++l) inc[w * 32 + l] = x[l];
            warp_tot[w] = x[31];
        }
        for (int t = 0; t < kScanThreads; ++t) {
            const int w = t >> 5, l = t & 31;
            const uint32_t k0 = uint32_t(t) * kScanItems;
            Run2 excl = l ? inc[t - 1] : Run2{0.0, 0.0};
            Run2 wpre{0.0, 0.0};
            for (int k = 0; k < w; ++k) wpre = run2_combine(wpre, warp_tot[k], double(32 * kScanItems));
            const Run2 et = run2_combine(wpre, excl, double(kScanItems * l));
            const Run2 full = run2_combine(tile_excl, et, double(k0));
            for (int j = 0; j < kScanItems; ++j) {
                if (k0 + j >= m) break;
                const uint64_t i = tile * uint64_t(kScanTile) + k0 + j;
                Pout[i] = full.p + LP[t][j];
                Qout[i] = fma(double(j + 1), full.p, full.q) + LQ[t][j];
                if (Sout) Sout[i] = V[t][j];
            }
        }
    }
};

template <class F, class Fn>
void parallel_tiles(const ScanPlan& P, int threads, Fn fn) {
    std::atomic<uint64_t> next(P.tile_lo);
    auto body = [&]() {
        for (;;) {
            const uint64_t t = next.fetch_add(1);
            if (t >= P.tile_hi) break;
            fn(t);
        }
    };
    if (threads <= 1) {
        body();
    } else {
        std::vector<std::thread> pool;
        for (int i = 0; i < threads; ++i) pool.emplace_back(body);
        for (auto& th : pool) th.join();
    }
}

// Phase 1 of a rank: aggregates and prefix tree of its own segments; totals[s][4] = (P_hi, P_lo,
// Q_hi, Q_lo) of every owned segment (what the kernel's segment duty stores into the comm blocks).
template <class F>
int run_scan_totals(const ScanPlan& P, const double* table, int threads, double* totals) {
    CpuScan<F> S(P, table);
    parallel_tiles<F>(P, threads, [&](uint64_t t) { S.tile_aggregate(t); });
    for (uint32_t s = P.seg_lo; s < P.seg_hi; ++s) {
        S.segment_prefixes(int(s));
        const Run& r = S.seg_tot[s];
        totals[4 * s + 0] = r.p.hi, totals[4 * s + 1] = r.p.lo, totals[4 * s + 2] = r.q.hi, totals[4 * s + 3] = r.q.lo;
    }
    return BQ_OK;
}

// Phase 2: with every segment's total known (all ranks'), the bases and this rank's slice of the
// tables.  Pout / Qout / Sout are indexed globally (the caller passes slice pointers shifted back).
template <class F>
int run_scan_finish(const ScanPlan& P, const double* table, int threads, const double* totals, double* Pout,
                    double* Qout, double* Sout, bq_scan_result* out) {
    CpuScan<F> S(P, table);
    parallel_tiles<F>(P, threads, [&](uint64_t t) { S.tile_aggregate(t); });
    for (uint32_t s = P.seg_lo; s < P.seg_hi; ++s) S.segment_prefixes(int(s));
    for (int s = 0; s < kSegments; ++s) {
        S.seg_tot[s].p = dd{totals[4 * s + 0], totals[4 * s + 1]};
        S.seg_tot[s].q = dd{totals[4 * s + 2], totals[4 * s + 3]};
    }
    S.bases();
    parallel_tiles<F>(P, threads, [&](uint64_t t) { S.tile_store(t, Pout, Qout, Sout); });
    memset(out, 0, sizeof(*out));
    out->slice_begin = P.tile_lo * uint64_t(kScanTile);
    const uint64_t e = P.tile_hi * uint64_t(kScanTile);
    out->slice_end = e < P.n ? e : P.n;
    if (out->slice_end == P.n && out->slice_end > out->slice_begin) {
        out->p_last = Pout[P.n - 1];
        out->q_last = Qout[P.n - 1];
        out->p_second_last = (P.n >= 2 && P.n - 2 >= out->slice_begin) ? Pout[P.n - 2] : 0.0;
    }
    out->world = P.world;
    out->rank = P.rank;
    out->status = BQ_OK;
    return BQ_OK;
}

template <class F>
int run_scan(const ScanPlan& P, const double* table, int threads, double* Pout, double* Qout,
             double* Sout, bq_scan_result* out) {
    double totals[4 * kSegments] = {};
    int rc = run_scan_totals<F>(P, table, threads, totals);
    if (rc != BQ_OK) return rc;
    rc = run_scan_finish<F>(P, table, threads, totals, Pout, Qout, Sout, out);
    if (rc == BQ_OK && P.n >= 2) out->p_second_last = Pout[P.n - 2];
    return rc;
}

template <class Fn>
int dispatch_scan(int integrand, Fn fn) {
    switch (integrand) {
        case BQ_F_SIN: return fn(SinF{});
        case BQ_F_ACC: return fn(AccF{});
        case BQ_F_VEL: return fn(VelF{});
        case BQ_F_DIS: return fn(DisF{});
        case BQ_F_PROFILE: return fn(ProfileF{});
        default: return BQ_ERR_INVALID;
    }
}

}  // namespace
}  // namespace bq

extern "C" int bq_cpu_cumulative(const bq_scan_job* job, int threads, double* P, double* Q,
                                 double* samples, bq_scan_result* out) {
    using namespace bq;
    if (!job || !P || !Q || !out) return BQ_ERR_INVALID;
    int len = 0;
    const double* table = default_profile(&len);
    ScanPlan plan;
    const int rc = make_scan_plan(*job, 0, 1, len, &plan);
    if (rc != BQ_OK) return rc;
    switch (plan.integrand) {
        case BQ_F_SIN: return run_scan<SinF>(plan, table, threads, P, Q, samples, out);
        case BQ_F_ACC: return run_scan<AccF>(plan, table, threads, P, Q, samples, out);
        case BQ_F_VEL: return run_scan<VelF>(plan, table, threads, P, Q, samples, out);
        case BQ_F_DIS: return run_scan<DisF>(plan, table, threads, P, Q, samples, out);
        case BQ_F_PROFILE: return run_scan<ProfileF>(plan, table, threads, P, Q, samples, out);
        default: return BQ_ERR_INVALID;
    }
}

// ---- one rank of a multi-process CPU world (the gloo layout; capability of the per-rank scans +
// root carry fix-up of 4main.c:90-160,178-221, without a root) ----------------------------------
extern "C" int bq_cpu_scan_totals(const bq_scan_job* job, int rank, int world, int threads, double* totals) {
    using namespace bq;
    if (!job || !totals) return BQ_ERR_INVALID;
    int len = 0;
    const double* table = default_profile(&len);
    ScanPlan plan;
    const int rc = make_scan_plan(*job, rank, world, len, &plan);
    if (rc != BQ_OK) return rc;
    return dispatch_scan(plan.integrand, [&](auto f) { return run_scan_totals<decltype(f)>(plan, table, threads, totals); });
}

extern "C" int bq_cpu_scan_finish(const bq_scan_job* job, int rank, int world, int threads, const double* totals,
                                  double* P_slice, double* Q_slice, double* samples_slice, bq_scan_result* out) {
    using namespace bq;
    if (!job || !totals || !P_slice || !Q_slice || !out) return BQ_ERR_INVALID;
    int len = 0;
    const double* table = default_profile(&len);
    ScanPlan plan;
    const int rc = make_scan_plan(*job, rank, world, len, &plan);
    if (rc != BQ_OK) return rc;
    const uint64_t shift = plan.tile_lo * uint64_t(kScanTile);     // the slice buffers start at this rank's first sample
    return dispatch_scan(plan.integrand, [&](auto f) {
        return run_scan_finish<decltype(f)>(plan, table, threads, totals, P_slice - shift, Q_slice - shift,
                                            samples_slice ? samples_slice - shift : nullptr, out);
    });
}
