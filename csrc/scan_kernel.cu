// The fused cumulative-integration kernel: ONE launch per GPU per call.
//
// Reference capability (4main.c:76-225): fill InterpProfile, per-rank running
// sum, gather to root, serial carry fix-up on the root, MPI_Bcast of 144 MB —
// twice.  Here, per GPU, in a single persistent launch:
//
//   phase A   every CTA reduces its tiles to (sum v, sum of local running sums);
//             f is evaluated on the fly, no sample array is ever materialised
//   phase A'  the last CTA to finish turns the tile aggregates of each owned
//             segment into exclusive prefixes (fixed-shape block scan, dd),
//             publishes the segment totals into every peer's comm block over
//             NVLink, meets the peers at an in-kernel barrier and folds the 8
//             totals in order into segment bases        (replaces 4main.c:134-155)
//   phase B   every CTA re-evaluates its tiles, scans them with the right base
//             and streams P and Q out — into every peer's table as well when
//             replicate is on (the fused all-gather that replaces MPI_Bcast,
//             4main.c:157,221)
//   end       second in-kernel barrier: when the kernel retires, every rank's
//             stores into this GPU's tables and the three final scalars are visible
//
// CTAs of one launch wait on each other, so the grid never exceeds what is
// co-resident (cooperative launch attribute).
#include <cuda_runtime.h>

#include <type_traits>

#include "ptx.cuh"
#include "quad_core.cuh"
#include "scan_core.cuh"
#include "scan_launch.h"

namespace bq {

namespace {

constexpr int kWarps = kScanThreads / 32;

struct ScanSmem {
    ScanPlan P;
    double red[2][kWarps];
    Run warp_tot[kWarps];
    Run2 warp_tot2[kWarps];
    Run seg_tot[kSegments];
    Run2 tile_excl;
    uint64_t gen;
    int is_last;
    int status;
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

__device__ __forceinline__ Run shfl_up_run(const Run& r, int d) {
    Run o;
    o.p.hi = __shfl_up_sync(0xffffffffu, r.p.hi, d);
    o.p.lo = __shfl_up_sync(0xffffffffu, r.p.lo, d);
    o.q.hi = __shfl_up_sync(0xffffffffu, r.q.hi, d);
    o.q.lo = __shfl_up_sync(0xffffffffu, r.q.lo, d);
    o.cnt = __shfl_up_sync(0xffffffffu, r.cnt, d);
    return o;
}

__device__ __forceinline__ bool spin_until(const uint64_t* flag, uint64_t target, bool sys,
                                           uint64_t limit_ns) {
    const uint64_t t0 = ptx::globaltimer();
    for (;;) {
        const uint64_t v = sys ? ptx::ld_acquire_sys(flag) : ptx::ld_acquire_gpu(flag);
        if (v >= target) return true;
        if (ptx::globaltimer() - t0 > limit_ns) return false;
        __nanosleep(40);
    }
}

// Exclusive scan of the tile aggregates of one segment by the whole CTA:
// thread t folds a contiguous block of tiles, a Kogge-Stone warp scan and a
// serial fold over the 8 warp totals stitch the blocks together.
__device__ void segment_prefixes(const ScanPlan& P, const ScanState& st, int seg, ScanSmem& sm) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t t0 = P.seg_tile[seg], t1 = P.seg_tile[seg + 1];
    const uint64_t T = t1 - t0;
    const uint64_t B = (T + kScanThreads - 1) / kScanThreads;
    const uint64_t lo = t0 + uint64_t(tid) * B < t1 ? t0 + uint64_t(tid) * B : t1;
    const uint64_t hi = lo + B < t1 ? lo + B : t1;

    Run mine = run_identity();
    for (uint64_t t = lo; t < hi; ++t) {
        const double* a = st.agg + 2 * (t - P.tile_lo);
        mine = run_combine(mine, run_from_tile(__ldcg(a), __ldcg(a + 1), double(scan_tile_count(P, t))));
    }
    Run inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const Run up = shfl_up_run(inc, d);
        if (lane >= d) inc = run_combine(up, inc);
    }
    if (lane == 31) sm.warp_tot[warp] = inc;
    __syncthreads();
    Run excl = shfl_up_run(inc, 1);
    if (lane == 0) excl = run_identity();
    Run wpre = run_identity();
    for (int w = 0; w < warp; ++w) wpre = run_combine(wpre, sm.warp_tot[w]);
    Run running = run_combine(wpre, excl);
    for (uint64_t t = lo; t < hi; ++t) {
        double* o = st.prefix + 4 * (t - P.tile_lo);
        o[0] = running.p.hi, o[1] = running.p.lo, o[2] = running.q.hi, o[3] = running.q.lo;
        const double* a = st.agg + 2 * (t - P.tile_lo);
        running = run_combine(running, run_from_tile(__ldcg(a), __ldcg(a + 1), double(scan_tile_count(P, t))));
    }
    if (tid == kScanThreads - 1) sm.seg_tot[seg] = running;
    __syncthreads();
}

template <class F>
__global__ void __launch_bounds__(kScanThreads) scan_kernel(ScanState st) {
    __shared__ ScanSmem sm;
    constexpr bool kIsProfile = std::is_same<F, ProfileF>::value;
    __shared__ __align__(16) double sTable[kIsProfile ? kMaxTable : 2];
    __shared__ __align__(8) uint64_t sBar;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (blockIdx.x == 0 && tid == 0) st.result->t_start_ns = ptx::globaltimer();
    {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(st.plan);
        uint32_t* dst = reinterpret_cast<uint32_t*>(&sm.P);
        for (int i = tid; i < int(sizeof(ScanPlan) / 4); i += kScanThreads) dst[i] = __ldg(src + i);
    }
    if (tid == 0) {
        sm.gen = ptx::ld_acquire_gpu(st.ready + 1) + 1;   // generation of this launch
        sm.status = BQ_OK;
        if (kIsProfile) ptx::mbar_init(&sBar, 1);
    }
    __syncthreads();
    const ScanPlan& P = sm.P;
    if (kIsProfile) {
        const uint32_t bytes = ((P.table_len + 1u) & ~1u) * 8u;
        if (tid == 0) {
            ptx::mbar_expect_tx(&sBar, bytes);
            ptx::tma_bulk_g2s(sTable, st.table, bytes, &sBar);
        }
        ptx::mbar_wait(&sBar, 0);
    }
    Plan fp;   // make_functor only reads p[] and table_len
    for (int i = 0; i < 4; ++i) fp.p[i] = P.p[i];
    fp.table_len = P.table_len;
    const F f = make_functor<F>(fp, sTable);
    const uint64_t limit = uint64_t(P.timeout_ms) * 1000000ull;
    const uint64_t gen = sm.gen;
    const bool multi = P.world > 1;

    // ---- phase A: tile aggregates -------------------------------------------
    for (uint64_t tile = P.tile_lo + blockIdx.x; tile < P.tile_hi; tile += gridDim.x) {
        const uint32_t m = scan_tile_count(P, tile);
        const uint32_t k0 = uint32_t(tid) * kScanItems;
        double v[kScanItems];
        scan_thread_values(f, P, tile * uint64_t(kScanTile) + k0, k0, m, v);
        double s, w;
        scan_thread_aggregate(v, k0, m, s, w);
        s = warp_sum(s);
        w = warp_sum(w);
        if (lane == 0) sm.red[0][warp] = s, sm.red[1][warp] = w;
        __syncthreads();
        if (warp == 0) {
            double a = lane < kWarps ? sm.red[0][lane] : 0.0;
            double b = lane < kWarps ? sm.red[1][lane] : 0.0;
#pragma unroll
            for (int off = 4; off >= 1; off >>= 1) {
                a += __shfl_xor_sync(0xffffffffu, a, off);
                b += __shfl_xor_sync(0xffffffffu, b, off);
            }
            if (lane == 0) {
                double* o = st.agg + 2 * (tile - P.tile_lo);
                o[0] = a, o[1] = b;
            }
        }
        __syncthreads();
    }
    if (tid == 0) {
        __threadfence();
        sm.is_last = (atomicAdd(st.ticket_a, 1u) == gridDim.x - 1);
    }
    __syncthreads();

    // ---- phase A': prefixes, cross-GPU exchange, segment bases -----------------
    if (sm.is_last) {
        __threadfence();
        if (tid < kSegments) sm.seg_tot[tid] = run_identity();
        __syncthreads();
        for (uint32_t s = P.seg_lo; s < P.seg_hi; ++s) segment_prefixes(P, st, int(s), sm);
        if (warp == 0) {
            int status = BQ_OK;
            if (multi) {
                const uint64_t epoch = st.comm->scan_epoch;
                const int par = int(epoch & 1);
                if (lane < P.world && !(P.flags & kScanSkipPublish)) {
                    CommBlock* peer = st.peers[lane];
                    for (uint32_t s = P.seg_lo; s < P.seg_hi; ++s) {
                        const Run r = sm.seg_tot[s];
                        ptx::st_relaxed_sys_f64x2(&peer->scan_slots[par][s][0], r.p.hi, r.p.lo);
                        ptx::st_relaxed_sys_f64x2(&peer->scan_slots[par][s][2], r.q.hi, r.q.lo);
                    }
                    ptx::red_release_sys_add(&peer->scan_arrived, 1ull);
                }
                __syncwarp();
                if (lane == 0 &&
                    !spin_until(&st.comm->scan_arrived, uint64_t(P.world) * (epoch + 1), true, limit))
                    status = BQ_ERR_TIMEOUT;
                status = __shfl_sync(0xffffffffu, status, 0);
                if (lane < kSegments) {
                    Run r;
                    ptx::ld_relaxed_sys_f64x2(&st.comm->scan_slots[par][lane][0], r.p.hi, r.p.lo);
                    ptx::ld_relaxed_sys_f64x2(&st.comm->scan_slots[par][lane][2], r.q.hi, r.q.lo);
                    sm.seg_tot[lane] = r;
                }
                __syncwarp();
            }
            if (lane == 0) {
                Run base = run_identity();
                for (int s = 0; s < kSegments; ++s) {
                    double* o = st.segbase + 4 * s;
                    o[0] = base.p.hi, o[1] = base.p.lo, o[2] = base.q.hi, o[3] = base.q.lo;
                    Run t = sm.seg_tot[s];
                    const uint64_t first = P.seg_tile[s] * uint64_t(kScanTile);
                    uint64_t last = P.seg_tile[s + 1] * uint64_t(kScanTile);
                    if (last > P.n) last = P.n;
                    t.cnt = double(last - first);
                    base = run_combine(base, t);
                }
                st.result->status = status;
                *st.ticket_a = 0u;
                __threadfence();
                ptx::st_release_gpu(st.ready, gen);
            }
        }
    }
    if (tid == 0) {
        if (!spin_until(st.ready, gen, false, 4 * limit)) sm.status = BQ_ERR_TIMEOUT;
    }
    __syncthreads();

    // ---- phase B: scan + store ------------------------------------------------
    const bool replicate = multi && (P.flags & kScanReplicate);
    for (uint64_t tile = P.tile_lo + blockIdx.x; tile < P.tile_hi; tile += gridDim.x) {
        const uint32_t m = scan_tile_count(P, tile);
        const uint32_t k0 = uint32_t(tid) * kScanItems;
        const uint64_t idx0 = tile * uint64_t(kScanTile) + k0;
        if (tid == 0) {
            const int seg = scan_segment_of(P, tile);
            const double* b = st.segbase + 4 * seg;
            const double* p = st.prefix + 4 * (tile - P.tile_lo);
            Run base{dd{__ldcg(b), __ldcg(b + 1)}, dd{__ldcg(b + 2), __ldcg(b + 3)}, 0.0};
            Run pre{dd{__ldcg(p), __ldcg(p + 1)}, dd{__ldcg(p + 2), __ldcg(p + 3)},
                    double((tile - P.seg_tile[seg]) * uint64_t(kScanTile))};
            sm.tile_excl = scan_tile_exclusive(base, pre);
        }
        double v[kScanItems], lp[kScanItems], lq[kScanItems];
        scan_thread_values(f, P, idx0, k0, m, v);
        scan_thread_local(v, lp, lq);
        Run2 inc{lp[kScanItems - 1], lq[kScanItems - 1]};
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            Run2 up{__shfl_up_sync(0xffffffffu, inc.p, d), __shfl_up_sync(0xffffffffu, inc.q, d)};
            if (lane >= d) inc = run2_combine(up, inc, double(kScanItems * d));
        }
        if (lane == 31) sm.warp_tot2[warp] = inc;
        Run2 excl{__shfl_up_sync(0xffffffffu, inc.p, 1), __shfl_up_sync(0xffffffffu, inc.q, 1)};
        if (lane == 0) excl = Run2{0.0, 0.0};
        __syncthreads();
        Run2 wpre{0.0, 0.0};
        for (int w = 0; w < warp; ++w) wpre = run2_combine(wpre, sm.warp_tot2[w], double(32 * kScanItems));
        const Run2 et = run2_combine(wpre, excl, double(kScanItems * lane));
        const Run2 full = run2_combine(sm.tile_excl, et, double(k0));
        double Pv[kScanItems], Qv[kScanItems];
#pragma unroll
        for (int j = 0; j < kScanItems; ++j) {
            Pv[j] = full.p + lp[j];
            Qv[j] = fma(double(j + 1), full.p, full.q) + lq[j];
        }
        if (k0 + kScanItems <= m) {
            ptx::st_f64x4(st.P + idx0, Pv[0], Pv[1], Pv[2], Pv[3]);
            ptx::st_f64x4(st.P + idx0 + 4, Pv[4], Pv[5], Pv[6], Pv[7]);
            ptx::st_f64x4(st.Q + idx0, Qv[0], Qv[1], Qv[2], Qv[3]);
            ptx::st_f64x4(st.Q + idx0 + 4, Qv[4], Qv[5], Qv[6], Qv[7]);
            if (st.samples) {
                ptx::st_f64x4(st.samples + idx0, v[0], v[1], v[2], v[3]);
                ptx::st_f64x4(st.samples + idx0 + 4, v[4], v[5], v[6], v[7]);
            }
            if (replicate) {
                for (int r = 0; r < P.world; ++r) {
                    if (r == P.rank) continue;
                    ptx::st_f64x4(st.peerP[r] + idx0, Pv[0], Pv[1], Pv[2], Pv[3]);
                    ptx::st_f64x4(st.peerP[r] + idx0 + 4, Pv[4], Pv[5], Pv[6], Pv[7]);
                    ptx::st_f64x4(st.peerQ[r] + idx0, Qv[0], Qv[1], Qv[2], Qv[3]);
                    ptx::st_f64x4(st.peerQ[r] + idx0 + 4, Qv[4], Qv[5], Qv[6], Qv[7]);
                }
            }
        } else {
#pragma unroll
            for (int j = 0; j < kScanItems; ++j) {
                if (k0 + j < m) {
                    st.P[idx0 + j] = Pv[j];
                    st.Q[idx0 + j] = Qv[j];
                    if (st.samples) st.samples[idx0 + j] = v[j];
                    if (replicate)
                        for (int r = 0; r < P.world; ++r)
                            if (r != P.rank) st.peerP[r][idx0 + j] = Pv[j], st.peerQ[r][idx0 + j] = Qv[j];
                }
            }
        }
        // The three scalars the callers print (4main.c:241 uses P[n-2]).
#pragma unroll
        for (int j = 0; j < kScanItems; ++j) {
            const uint64_t i = idx0 + j;
            if (i + 1 == P.n || i + 2 == P.n) {
                for (int r = 0; r < P.world; ++r) {
                    double* fin = st.peers[r]->scan_final;
                    if (i + 1 == P.n) {
                        ptx::st_relaxed_sys_f64(fin + 0, Pv[j]);
                        ptx::st_relaxed_sys_f64(fin + 1, Qv[j]);
                        if (P.n == 1) ptx::st_relaxed_sys_f64(fin + 2, 0.0);
                    } else {
                        ptx::st_relaxed_sys_f64(fin + 2, Pv[j]);
                    }
                }
            }
        }
        __syncthreads();
    }

    // ---- end: second barrier, result -------------------------------------------
    __syncthreads();
    if (tid == 0) {
        if (multi) ptx::fence_sys(); else __threadfence();
        sm.is_last = (atomicAdd(st.ticket_b, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!sm.is_last || warp != 0) return;
    int status = sm.status;
    if (multi) {
        ptx::fence_sys();
        const uint64_t epoch = st.comm->scan_epoch;
        if (lane < P.world && !(P.flags & kScanSkipPublish))
            ptx::red_release_sys_add(&st.peers[lane]->scan_done, 1ull);
        __syncwarp();
        if (lane == 0) {
            if (!spin_until(&st.comm->scan_done, uint64_t(P.world) * (epoch + 1), true, limit))
                status = BQ_ERR_TIMEOUT;
            st.comm->scan_epoch = epoch + 1;
        }
    }
    if (lane == 0) {
        bq_scan_result* R = st.result;
        R->p_last = ptx::ld_relaxed_sys_f64(st.comm->scan_final + 0);
        R->q_last = ptx::ld_relaxed_sys_f64(st.comm->scan_final + 1);
        R->p_second_last = ptx::ld_relaxed_sys_f64(st.comm->scan_final + 2);
        R->slice_begin = P.tile_lo * uint64_t(kScanTile);
        uint64_t e = P.tile_hi * uint64_t(kScanTile);
        R->slice_end = e < P.n ? e : P.n;
        if (status != BQ_OK) R->status = status;
        R->world = P.world;
        R->rank = P.rank;
        *st.ticket_b = 0u;
        st.ready[1] = gen;             // generation completed
        R->t_end_ns = ptx::globaltimer();
    }
}

template <class F>
cudaError_t launch_scan_f(const ScanState& st, unsigned grid, cudaStream_t stream) {
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(grid, 1, 1);
    lc.blockDim = dim3(kScanThreads, 1, 1);
    lc.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    lc.attrs = attr;
    lc.numAttrs = 1;
    return cudaLaunchKernelEx(&lc, scan_kernel<F>, st);
}

template <class F>
int occupancy_f() {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, scan_kernel<F>, kScanThreads, 0) != cudaSuccess) {
        cudaGetLastError();
        return 1;
    }
    return n < 1 ? 1 : n;
}

}  // namespace

int scan_max_grid(int /*device*/, int sm_count) {
    int occ = occupancy_f<ProfileF>();
    const int o2 = occupancy_f<DisF>();
    if (o2 < occ) occ = o2;
    if (occ > 4) occ = 4;
    return occ * sm_count;
}

cudaError_t launch_scan(const ScanPlan& P, const ScanState& st, unsigned grid, cudaStream_t stream) {
    switch (P.integrand) {
        case BQ_F_SIN: return launch_scan_f<SinF>(st, grid, stream);
        case BQ_F_ACC: return launch_scan_f<AccF>(st, grid, stream);
        case BQ_F_VEL: return launch_scan_f<VelF>(st, grid, stream);
        case BQ_F_DIS: return launch_scan_f<DisF>(st, grid, stream);
        case BQ_F_PROFILE: return launch_scan_f<ProfileF>(st, grid, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace bq
