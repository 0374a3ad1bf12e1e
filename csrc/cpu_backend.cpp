// CPU backend: walks exactly the summation tree of the sm_100a kernel
// (quad_kernel.cu) with std::thread workers, so a box without a GPU can run
// the API, the CLIs and the tests.  It plays the role the MPI programs play in
// the reference (riemann.cpp:47-101, 4main.c:42-247) but with every "rank"
// being a chunk of the fixed tree instead of a slice that depends on -n.
#include <math.h>
#include <string.h>

#include <atomic>
#include <thread>
#include <type_traits>
#include <vector>

#include "host_common.h"
#include "quad_core.cuh"

namespace bq {

static inline void host_sincos(double x, double* s, double* c) { ::sincos(x, s, c); }

// The 32-lane xor butterfly of warp_reduce_dd(); lane 0's value.
static dd butterfly(dd* v, int lanes) {
    dd tmp[32];
    for (int off = lanes / 2; off >= 1; off >>= 1) {
        for (int l = 0; l < lanes; ++l) tmp[l] = dd_add(v[l], v[l ^ off]);
        for (int l = 0; l < lanes; ++l) v[l] = tmp[l];
    }
    return v[0];
}

// The xor butterfly over plain doubles (the kernel's warp sum of the diffusion residues).
static double butterfly_sum(double* v, int lanes) {
    double tmp[32];
    for (int off = lanes / 2; off >= 1; off >>= 1) {
        for (int l = 0; l < lanes; ++l) tmp[l] = v[l] + v[l ^ off];
        for (int l = 0; l < lanes; ++l) v[l] = tmp[l];
    }
    return v[0];
}

// warp_reduce_lazy(): same shape, un-normalised pairs.
static dd butterfly_lazy(dd* v, int lanes) {
    dd tmp[32];
    for (int off = lanes / 2; off >= 1; off >>= 1) {
        for (int l = 0; l < lanes; ++l) tmp[l] = dd_add_lazy(v[l], v[l ^ off]);
        for (int l = 0; l < lanes; ++l) v[l] = tmp[l];
    }
    return v[0];
}

// PartWorker::tile_rotate, fused fp32: the ragged last row's share of the column constants' rounding.
static inline void ragged_row_owed(TileAcc<float>& acc, int j, float S, float C, double plo, double qlo) {
    acc.d[j] += fma(double(S), plo, double(C) * qlo);
}
static inline void ragged_row_owed(TileAcc<double>&, int, double, double, double, double) {}

template <class F, class R>
struct CpuWorker {
    const Plan& P;
    F f;
    bool rotate;
    // column constants per virtual thread, row rotations per virtual lane
    std::vector<R> cj, sj;
    std::vector<double> dx, plo_col, qlo_col;   // per column: constant - its fp32 value (fused fp32, ragged last row)
    double owed4[4] = {0.0, 0.0, 0.0, 0.0};   // sums over the 2048 columns: float(P), float(Q), P - float(P), Q - float(Q)
    SinCos rowrot[32];

    static SinCos host_sc(double ang) {
        SinCos r;
        host_sincos(ang, &r.s, &r.c);
        return r;
    }

    CpuWorker(const Plan& P_, const double* table)
        : P(P_), f(make_functor<F>(P_, table)), rotate(P_.eval == BQ_EVAL_ROTATE) {
        cj.resize(kThreads * kCols);
        sj.resize(kThreads * kCols);
        dx.resize(kThreads * kCols);
        plo_col.assign(kThreads * kCols, 0.0);
        qlo_col.assign(kThreads * kCols, 0.0);
        std::vector<double> psum(kThreads, 0.0), qsum(kThreads, 0.0), plo(kThreads, 0.0), qlo(kThreads, 0.0);
        SinCos blk[kCols];
        for (int j = 0; j < kCols; ++j) blk[j] = host_sc(f.angle(double(j * kThreads) * P.h));
        for (int t = 0; t < kThreads; ++t) {
            const SinCos mine = host_sc(f.angle(double(t) * P.h));
            for (int j = 0; j < kCols; ++j) {
                const int idx = t * kCols + j;
                if (rotate) {
                    const SinCos cs = rotate_sc(mine, blk[j]);
                    if (F::kFusedSum) {
                        double Pc, Qc;
                        fused_consts(f, fused_norm<F, R>(f), cs, Pc, Qc);
                        cj[idx] = fused_hi<R>(Pc, (P.flags & kFlagFineGrid) != 0u);
                        sj[idx] = R(Qc);
                        const double ph = double(cj[idx]), qh = double(sj[idx]);
                        psum[t] += ph, qsum[t] += qh;
                        plo[t] += Pc - ph, qlo[t] += Qc - qh;
                        plo_col[idx] = Pc - ph, qlo_col[idx] = Qc - qh;
                    } else {
                        cj[idx] = R(cs.c);
                        sj[idx] = R(cs.s);
                    }
                    dx[idx] = double(t + kThreads * j) * P.h;
                } else {
                    cj[idx] = R(0);
                    sj[idx] = R(0);
                    dx[idx] = double(t + kThreads * j);
                }
            }
        }
        for (int l = 0; l < 32; ++l) rowrot[l] = host_sc(f.angle(double(l) * double(kWidth) * P.h));
        // warp butterflies, then the 8 warp sums in order (PartWorker's constructor)
        std::vector<double>* parts4[4] = {&psum, &qsum, &plo, &qlo};
        for (int i = 0; i < 4; ++i)
            for (int w = 0; w < kThreads / 32; ++w) owed4[i] += butterfly_sum(parts4[i]->data() + w * 32, 32);
    }

    dd part(uint32_t part_idx) const {
        uint64_t lo, hi;
        part_tiles(P, part_idx, lo, hi);
        dd total[kThreads];
        TileAcc<R> accs[kThreads];
        for (int t = 0; t < kThreads; ++t) total[t] = dd{0.0, 0.0}, accs[t].clear();
        R S[kMaxRows], C[kMaxRows];
        double X0[kMaxRows];
        RowDiffusion<R, F::kFusedSum> diff[kMaxRows];     // per row lane, carried through the part
        constexpr bool uniform_rows = F::kFusedSum && sizeof(R) == 8;   // PartWorker::kUniformRows
        RowConsts rowc;
        if (rotate && uniform_rows) fill_row_consts(P, &rowc);
        for (uint64_t tile = lo; tile < hi; ++tile) {
            if (rotate && uniform_rows && ((tile - lo) & (kFlushTiles - 1)) == 0) {
                // PartWorker::run_uniform: a flush group of full tiles, none of them the grid's last
                const uint64_t nt = hi - tile < uint64_t(kFlushTiles) ? hi - tile : uint64_t(kFlushTiles);
                if (tile + nt - 1 < P.n_tiles - 1) {
                    const SinCos tb = host_sc(f.angle(fma(double(tile * P.tile_pts) + P.off, P.h, P.a)));
                    const bool fine = (P.flags & kFlagFineGrid) != 0u;
                    for (int t = 0; t < kThreads; ++t) {
                        R A[kCols], B[kCols];
                        for (int j = 0; j < kCols; ++j) {
                            const R pj = cj[t * kCols + j], qj = sj[t * kCols + j];
                            A[j] = R(fma(tb.s, double(pj), tb.c * double(qj)));
                            B[j] = R(fma(tb.c, double(pj), -(tb.s * double(qj))));
                        }
                        TileAcc<R>& acc = accs[t];
                        for (uint64_t q = 0; q < nt * P.rows; ++q)      // nt * rows consecutive rows
                            for (int j = 0; j < kCols; ++j)
                                acc.add_rot(j, R(rowc.cs[q][0]), R(rowc.cs[q][1]), A[j], B[j]);
                        total[t] = dd_acc_lazy(total[t], fused_norm<F, R>(f).amp * acc.flush_fused(fine));
                        acc.clear();
                    }
                    tile += nt - 1;
                    continue;
                }
            }
            const TileGeom g = tile_geom(P, tile);
            if (rotate) {
                // every warp of the kernel derives the same row bases for itself
                const SinCos tb = host_sc(f.angle(fma(double(tile * P.tile_pts) + P.off, P.h, P.a)));
                for (uint32_t k = 0; k < kMaxRows; ++k) {
                    SinCos rb = rotate_sc(tb, rowrot[k]);
                    if (F::kFusedSum && fused_norm<F, R>(f).swap) rb = SinCos{rb.c, rb.s};
                    diff[k].round(rb, k < g.full_rows, S[k], C[k]);     // full rows: every column uses them
                    X0[k] = fma(row_base_index(P, g, k), P.h, P.a);
                }
            }
            for (int t = 0; t < kThreads; ++t) {
                const R* cjt = &cj[t * kCols];
                const R* sjt = &sj[t * kCols];
                const double* dxt = &dx[t * kCols];
                TileAcc<R>& acc = accs[t];
                double rb = row_base_index(P, g, 0);
                constexpr bool fused = F::kFusedSum;
                const bool fine = (P.flags & kFlagFineGrid) != 0u;
                const int fast = !(F::kTrig && (P.flags & kFlagFastTrig)) ? 0 : ((P.flags & kFlagRowK) ? 2 : 1);
                double dcol[kCols];
                for (int j = 0; j < kCols; ++j) dcol[j] = dxt[j];
                for (uint32_t r = 0; r < g.full_rows; ++r) {
                    if (rotate && fused) {
                        for (int j = 0; j < kCols; ++j) acc.add_rot(j, S[r], C[r], cjt[j], sjt[j]);
                        if ((r & (kPromoteRows - 1)) == kPromoteRows - 1) acc.promote(fine);
                    } else {
                        R v[kCols];
                        if (rotate) {
                            for (int j = 0; j < kCols; ++j)
                                v[j] = point_rotate<F, R>(f, S[r], C[r], cjt[j], sjt[j], X0[r], dxt[j]);
                        } else if (fast == 2) {
                            RowDirect<F, R, 2>::eval(f, rb, dcol, P.h, P.a, v);
                        } else if (fast == 1) {
                            RowDirect<F, R, 1>::eval(f, rb, dcol, P.h, P.a, v);
                        } else {
                            RowDirect<F, R, 0>::eval(f, rb, dcol, P.h, P.a, v);
                        }
                        acc.add_row(v);
                    }
                    rb += double(kWidth);
                }
                if (g.rem) {
                    const uint32_t r = g.full_rows;
                    for (int j = 0; j < kCols; ++j)
                        if (uint32_t(t + kThreads * j) < g.rem) {
                            if (rotate && fused) {
                                acc.add_rot(j, S[r], C[r], cjt[j], sjt[j]);
                                ragged_row_owed(acc, j, S[r], C[r], plo_col[t * kCols + j], qlo_col[t * kCols + j]);
                            } else
                                acc.add_one(j, rotate ? point_rotate<F, R>(f, S[r], C[r], cjt[j], sjt[j],
                                                                           X0[r], dxt[j])
                                                      : fast ? point_direct<F, R, 1>(f, rb, dxt[j], P.h, P.a)
                                                             : point_direct<F, R, 0>(f, rb, dxt[j], P.h, P.a));
                        }
                }
                if (((tile - lo) & (kFlushTiles - 1)) == kFlushTiles - 1 || tile + 1 == hi) {
                    total[t] = dd_acc_lazy(total[t], (rotate && fused) ? fused_norm<F, R>(f).amp * acc.flush_fused(fine) : acc.flush());
                    acc.clear();
                }
            }
        }
        if (rotate && RowDiffusion<R, F::kFusedSum>::kOn) {
            // what the row values still owe the sums (RowDiffusion)
            for (int k = 0; k < kMaxRows; ++k)      // lane k of the first warp settles row stream k
                total[k] = dd_acc_lazy(total[k], fused_norm<F, R>(f).amp * diffusion_owed(diff[k].es, diff[k].ec, diff[k].ss,
                                                                                        diff[k].sc, owed4[0], owed4[1],
                                                                                        owed4[2], owed4[3]));
        }
        // block_reduce_dd(): 8 warp butterflies, then an 8-lane butterfly, one renormalisation.
        dd wsum[8];
        for (int w = 0; w < kThreads / 32; ++w) wsum[w] = butterfly_lazy(&total[w * 32], 32);
        return dd_renorm(butterfly_lazy(wsum, 8));
    }

    dd chunk(uint32_t c) const {
        const dd p0 = part(c * kPartsPerChunk + 0);
        const dd p1 = part(c * kPartsPerChunk + 1);
        return dd_add(p0, p1);
    }
};

template <>
inline CustomHostF make_functor<CustomHostF>(const Plan& P, const double*) {
    return CustomHostF{custom_eval(P.integrand), {P.p[0], P.p[1], P.p[2], P.p[3]}};
}

static dd segment_sum(const dd* partials, uint32_t seg) {
    dd lanes[32];
    const uint32_t c0 = seg * kChunksPerSegment;
    for (int l = 0; l < 32; ++l) {
        dd acc{0.0, 0.0};
        for (uint32_t c = c0 + l; c < c0 + kChunksPerSegment; c += 32) acc = dd_add(acc, partials[c]);
        lanes[l] = acc;
    }
    return butterfly(lanes, 32);
}

template <class F, class R>
static void run_segments(const Plan& P, const double* table, int threads, double* seg_hi,
                         double* seg_lo) {
    CpuWorker<F, R> worker(P, table);
    std::vector<dd> partials(kChunks, dd{0.0, 0.0});
    std::atomic<uint32_t> next(P.chunk_lo);
    auto body = [&]() {
        for (;;) {
            const uint32_t c = next.fetch_add(1);
            if (c >= P.chunk_hi) break;
            partials[c] = worker.chunk(c);
        }
    };
    if (threads <= 1) {
        body();
    } else {
        std::vector<std::thread> pool;
        for (int i = 0; i < threads; ++i) pool.emplace_back(body);
        for (auto& th : pool) th.join();
    }
    for (uint32_t s = P.seg_lo; s < P.seg_hi; ++s) {
        const dd v = segment_sum(partials.data(), s);
        seg_hi[s] = v.hi;
        seg_lo[s] = v.lo;
    }
}

template <class F>
static void run_f(const Plan& P, const double* table, int threads, double* hi, double* lo) {
    if (P.dtype == BQ_F64)
        run_segments<F, double>(P, table, threads, hi, lo);
    else
        run_segments<F, float>(P, table, threads, hi, lo);
}

static int cpu_segments(const bq_job* job, int rank, int world, int threads, const double* table,
                        int table_len, Plan* P, double* hi, double* lo) {
    int table_default_len = 0;
    if (!table) table = default_profile(&table_default_len), table_len = table_default_len;
    const int rc = make_plan(*job, rank, world, 0, table_len, P);
    if (rc != BQ_OK) return rc;
    for (int s = 0; s < kSegments; ++s) hi[s] = lo[s] = 0.0;
    switch (P->integrand) {
        case BQ_F_SIN: run_f<SinF>(*P, table, threads, hi, lo); break;
        case BQ_F_ACC: run_f<AccF>(*P, table, threads, hi, lo); break;
        case BQ_F_VEL: run_f<VelF>(*P, table, threads, hi, lo); break;
        case BQ_F_DIS: run_f<DisF>(*P, table, threads, hi, lo); break;
        case BQ_F_PROFILE: run_f<ProfileF>(*P, table, threads, hi, lo); break;
        default:
            if (P->integrand < BQ_F_CUSTOM || !custom_eval(P->integrand)) {
                set_error("custom integrand has no host evaluator");
                return BQ_ERR_INVALID;
            }
            run_f<CustomHostF>(*P, table, threads, hi, lo);
            break;
    }
    return BQ_OK;
}

int fold_segments(const Plan& P, const double* table, const double* hi, const double* lo,
                  bq_result* out) {
    switch (P.integrand) {
        case BQ_F_SIN: finish_result(P, make_functor<SinF>(P, table), hi, lo, out); break;
        case BQ_F_ACC: finish_result(P, make_functor<AccF>(P, table), hi, lo, out); break;
        case BQ_F_VEL: finish_result(P, make_functor<VelF>(P, table), hi, lo, out); break;
        case BQ_F_DIS: finish_result(P, make_functor<DisF>(P, table), hi, lo, out); break;
        case BQ_F_PROFILE: finish_result(P, make_functor<ProfileF>(P, table), hi, lo, out); break;
        default:
            if (P.integrand < BQ_F_CUSTOM || !custom_eval(P.integrand)) return BQ_ERR_INVALID;
            finish_result(P, make_functor<CustomHostF>(P, table), hi, lo, out);
            break;
    }
    out->status = BQ_OK;
    return BQ_OK;
}

}  // namespace bq

extern "C" {

int bq_cpu_integrate_segments(const bq_job* job, int rank, int world, int threads, bq_result* out) {
    if (!job || !out) return BQ_ERR_INVALID;
    memset(out, 0, sizeof(*out));
    bq::Plan P;
    return bq::cpu_segments(job, rank, world, threads, nullptr, 0, &P, out->seg_hi, out->seg_lo);
}

int bq_cpu_integrate(const bq_job* job, int threads, bq_result* out) {
    if (!job || !out) return BQ_ERR_INVALID;
    memset(out, 0, sizeof(*out));
    bq::Plan P;
    double hi[BQ_SEGMENTS], lo[BQ_SEGMENTS];
    const int rc = bq::cpu_segments(job, 0, 1, threads, nullptr, 0, &P, hi, lo);
    if (rc != BQ_OK) return rc;
    int len = 0;
    return bq::fold_segments(P, bq::default_profile(&len), hi, lo, out);
}

int bq_fold_segments(const bq_job* job, const double* seg_hi, const double* seg_lo, bq_result* out) {
    if (!job || !out || !seg_hi || !seg_lo) return BQ_ERR_INVALID;
    bq::Plan P;
    int len = 0;
    const double* table = bq::default_profile(&len);
    const int rc = bq::make_plan(*job, 0, 1, 0, len, &P);
    if (rc != BQ_OK) return rc;
    double hi[BQ_SEGMENTS], lo[BQ_SEGMENTS];
    memcpy(hi, seg_hi, sizeof(hi));
    memcpy(lo, seg_lo, sizeof(lo));
    memset(out, 0, sizeof(*out));
    return bq::fold_segments(P, table, hi, lo, out);
}

}  // extern "C"
