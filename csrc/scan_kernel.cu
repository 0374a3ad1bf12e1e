// The fused cumulative-integration kernel: ONE launch per GPU per call.
//
// Reference capability (4main.c:76-225): fill InterpProfile, per-rank running
// sum, gather to root, serial carry fix-up on the root, MPI_Bcast of 144 MB —
// twice.  Here, per GPU, in a single persistent launch:
//
//   phase A   every WARP reduces whole tiles to (sum v, sum of local running
//             sums); f is evaluated on the fly, no sample array is ever
//             materialised, and no CTA-wide barrier sits in the tile loop
//   phase A'  once every tile aggregate is in, one CTA per owned segment turns
//             that segment's aggregates into exclusive prefixes (fixed-shape
//             block scan, dd) — the segments in parallel; the last of them
//             publishes the segment totals into every peer's comm block over
//             NVLink, meets the peers at an in-kernel barrier and folds the 8
//             totals in order into segment bases        (replaces 4main.c:134-155)
//   phase B   every warp re-evaluates its tiles, scans them with the right base
//             and streams P and Q out in 256-bit stores — into every peer's
//             table as well when replicate is on (the fused all-gather that
//             replaces MPI_Bcast, 4main.c:157,221)
//   end       second in-kernel barrier: when the kernel retires, every rank's
//             stores into this GPU's tables and the three final scalars are visible
//
// A tile is 2048 samples = 8 passes of 32 lanes x 8 samples.  The arithmetic
// tree of a tile (scan_core.cuh) is that of a 256-thread block scan; one warp
// walks it pass by pass, so the result does not depend on which warp or CTA
// ran the tile, and matches cpu_scan.cpp bit for bit.
//
// CTAs of one launch wait on each other, so the grid never exceeds what is
// co-resident (cooperative launch attribute).
#include <cuda_runtime.h>
#include <stdlib.h>

#include <atomic>
#include <type_traits>

#include "ptx.cuh"
#include "quad_core.cuh"
#include "scan_core.cuh"
#include "scan_launch.h"

namespace bq {

namespace {

constexpr int kWarps = kScanThreads / 32;
constexpr size_t kScanStageBytes = size_t(kWarps) * 2 * 32 * kScanItems * sizeof(double);   // 32 KB

struct ScanSmem {
    Run warp_tot[kWarps];
    Run seg_tot[kSegments];
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

    // Both sequential folds below read their tile aggregates in batches: the loads
    // of a batch are all in flight before the first dependent dd combine, instead
    // of one L2 round trip per tile on the critical path.
    constexpr int kBatch = 8;
    Run mine = run_identity();
    for (uint64_t t = lo; t < hi; t += kBatch) {
        double s[kBatch], w[kBatch];
#pragma unroll
        for (int k = 0; k < kBatch; ++k) {
            const uint64_t tt = t + k < hi ? t + k : hi - 1;
            const double* a = st.agg + 2 * (tt - P.tile_lo);
            s[k] = __ldcg(a), w[k] = __ldcg(a + 1);
        }
#pragma unroll
        for (int k = 0; k < kBatch; ++k)
            if (t + k < hi) mine = run_combine(mine, run_from_tile(s[k], w[k], double(scan_tile_count(P, t + k))));
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
    for (uint64_t t = lo; t < hi; t += kBatch) {
        double s[kBatch], w[kBatch];
#pragma unroll
        for (int k = 0; k < kBatch; ++k) {
            const uint64_t tt = t + k < hi ? t + k : hi - 1;
            const double* a = st.agg + 2 * (tt - P.tile_lo);
            s[k] = __ldcg(a), w[k] = __ldcg(a + 1);
        }
#pragma unroll
        for (int k = 0; k < kBatch; ++k) {
            if (t + k < hi) {
                double* o = st.prefix + 4 * (t + k - P.tile_lo);
                ptx::st_f64x4(o, running.p.hi, running.p.lo, running.q.hi, running.q.lo);
                running = run_combine(running, run_from_tile(s[k], w[k], double(scan_tile_count(P, t + k))));
            }
        }
    }
    if (tid == kScanThreads - 1) sm.seg_tot[seg] = running;
    __syncthreads();
}

struct ScanParams {
    ScanPlan P;        // rides in the kernel parameter buffer
    ScanState st;
};

// Two resident CTAs per SM (128 registers): measured best on B200 — the register room lets
// the compiler overlap the passes of a tile; 3 and 4 CTAs/SM (80 / 64 registers) both lost.
// REPL = the fused all-gather (replicate) is compiled in; a separate instantiation so the
// single-GPU / slice-only kernel carries none of its registers or staging.
template <class F, bool REPL>
__global__ void __launch_bounds__(kScanThreads, 2) scan_kernel(const __grid_constant__ ScanParams prm) {
    const ScanPlan& P = prm.P;
    const ScanState& st = prm.st;
    __shared__ ScanSmem sm;
    constexpr bool kIsProfile = std::is_same<F, ProfileF>::value;
    __shared__ __align__(16) double sTable[kIsProfile ? kMaxTable : 2];
    __shared__ __align__(8) uint64_t sBar;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (blockIdx.x == 0 && tid == 0) st.result->t_start_ns = ptx::globaltimer();
    if (tid == 0) {
        sm.gen = ptx::ld_acquire_gpu(st.ready + 1) + 1;   // generation of this launch
        sm.status = BQ_OK;
        if (kIsProfile) ptx::mbar_init(&sBar, 1);
    }
    __syncthreads();
    if (kIsProfile) {
        const uint32_t bytes = ((P.table_len + 1u) & ~1u) * 8u;
        if (tid == 0) {
            ptx::mbar_expect_tx(&sBar, bytes);
            ptx::tma_bulk_g2s(sTable, st.table, bytes, &sBar);
        }
        ptx::mbar_wait(&sBar, 0);
    }
    const F f = make_scan_functor<F>(P, sTable);
    const uint64_t limit = uint64_t(P.timeout_ms) * 1000000ull;
    const uint64_t gen = sm.gen;
    const bool multi = P.world > 1;
    const uint64_t gwarp = uint64_t(blockIdx.x) * kWarps + warp;
    const uint64_t nwarps = uint64_t(gridDim.x) * kWarps;

    // ---- phase A: tile aggregates, one warp per tile ---------------------------
    for (uint64_t tile = P.tile_lo + gwarp; tile < P.tile_hi; tile += nwarps) {
        const uint32_t m = scan_tile_count(P, tile);
        // pass j = the samples of "warp j" of a block scan; each lane folds its 8 pass
        // shares in the order ((0+4)+(2+6)) + ((1+5)+(3+7)), the 32 lanes then meet in
        // one xor butterfly (scan tree of cpu_scan.cpp: tile_aggregate)
        auto pass = [&](int j, double& s_out, double& w_out) {
            const uint32_t k0 = uint32_t(j * 32 + lane) * kScanItems;
            double v[kScanItems];
            scan_thread_values(f, P, tile * uint64_t(kScanTile) + k0, k0, m, v);
            scan_thread_aggregate(v, k0, m, s_out, w_out);   // this lane's share of the pass
        };
        auto pair = [&](int j, double& s_out, double& w_out) {
            double s0, w0, s1, w1;
            pass(j, s0, w0);
            pass(j + 4, s1, w1);
            s_out = s0 + s1;
            w_out = w0 + w1;
        };
        double sa, wa, sb, wb, sc, wc, sd, wd;
        pair(0, sa, wa);
        pair(2, sb, wb);
        sa += sb, wa += wb;
        pair(1, sc, wc);
        pair(3, sd, wd);
        sc += sd, wc += wd;
        // the lane's 8 pass shares are folded first, then ONE butterfly per tile
        const double S = warp_sum(sa + sc), W = warp_sum(wa + wc);
        if (lane == 0) ptx::st_f64x2(st.agg + 2 * (tile - P.tile_lo), S, W);
    }
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        atomicAdd(st.ticket_a, 1u);
    }

    // ---- phase A': prefixes (segments in parallel), exchange, segment bases -----
    const uint32_t owned = P.seg_hi - P.seg_lo;
    const uint32_t workers = owned < gridDim.x ? owned : gridDim.x;
    if (blockIdx.x < workers) {
        if (tid == 0) {
            const uint64_t t0 = ptx::globaltimer();
            while (ptx::ld_acquire_gpu_u32(st.ticket_a) < gridDim.x) {
                if (ptx::globaltimer() - t0 > 4 * limit) {
                    sm.status = BQ_ERR_TIMEOUT;
                    break;
                }
            }
        }
        if (tid < kSegments) sm.seg_tot[tid] = run_identity();
        __syncthreads();
        __threadfence();
        for (uint32_t s = P.seg_lo + blockIdx.x; s < P.seg_hi; s += workers) {
            segment_prefixes(P, st, int(s), sm);
            if (tid == 0) {
                // This segment's total goes where a peer would publish it: the slot of
                // this call's parity in the own comm block, for the CTA that folds the bases.
                const Run r = sm.seg_tot[s];
                double* o = &st.comm->scan_slots[int(st.comm->scan_epoch & 1)][s][0];
                ptx::st_relaxed_sys_f64x2(o, r.p.hi, r.p.lo);
                ptx::st_relaxed_sys_f64x2(o + 2, r.q.hi, r.q.lo);
            }
        }
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            sm.is_last = (atomicAdd(st.ticket_a + 1, 1u) == workers - 1);
        }
        __syncthreads();
        if (sm.is_last && warp == 0) {
            __threadfence();
            int status = sm.status;
            if (lane == 0) st.result->reserved[0] = ptx::globaltimer();   // all prefixes done
            if (lane < kSegments) {
                Run r = run_identity();
                if (uint32_t(lane) >= P.seg_lo && uint32_t(lane) < P.seg_hi) {
                    const double* o = &st.comm->scan_slots[int(st.comm->scan_epoch & 1)][lane][0];
                    ptx::ld_relaxed_sys_f64x2(o, r.p.hi, r.p.lo);
                    ptx::ld_relaxed_sys_f64x2(o + 2, r.q.hi, r.q.lo);
                }
                sm.seg_tot[lane] = r;
            }
            __syncwarp();
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
                    r.cnt = 0.0;
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
                st.ticket_a[0] = 0u;
                st.ticket_a[1] = 0u;
                st.result->reserved[1] = ptx::globaltimer();                // bases ready
                __threadfence();
                ptx::st_release_gpu(st.ready, gen);
            }
        }
    }
    if (tid == 0) {
        if (!spin_until(st.ready, gen, false, 4 * limit)) sm.status = BQ_ERR_TIMEOUT;
    }
    __syncthreads();

    // ---- phase B: scan + store, one warp per tile --------------------------------
    const bool replicate = REPL && multi && (P.flags & kScanReplicate);
    extern __shared__ __align__(32) double sStage[];   // replicate only: [warp][P 256 | Q 256]
    double* const stage = sStage + warp * (2 * 32 * kScanItems);
    for (uint64_t tile = P.tile_lo + gwarp; tile < P.tile_hi; tile += nwarps) {
        const uint32_t m = scan_tile_count(P, tile);
        // exclusive prefix of the tile: segment base (+) in-segment prefix
        Run2 tile_excl;
        {
            const int seg = scan_segment_of(P, tile);
            const double* bp = st.segbase + 4 * seg;
            const double* pp = st.prefix + 4 * (tile - P.tile_lo);
            const Run base{dd{__ldcg(bp), __ldcg(bp + 1)}, dd{__ldcg(bp + 2), __ldcg(bp + 3)}, 0.0};
            const Run pre{dd{__ldcg(pp), __ldcg(pp + 1)}, dd{__ldcg(pp + 2), __ldcg(pp + 3)},
                          double((tile - P.seg_tile[seg]) * uint64_t(kScanTile))};
            tile_excl = scan_tile_exclusive(base, pre);
        }
        Run2 wpre{0.0, 0.0};   // fold of the passes before this one ("warps before mine")
#pragma unroll 1
        for (int j = 0; j < kWarps; ++j) {
            const uint32_t k0 = uint32_t(j * 32 + lane) * kScanItems;
            const uint64_t idx0 = tile * uint64_t(kScanTile) + k0;
            double v[kScanItems];
            scan_thread_values(f, P, idx0, k0, m, v);
            Run2 inc;
            {
                double p = 0.0, q = 0.0;
#pragma unroll
                for (int i = 0; i < kScanItems; ++i) p += v[i], q += p;
                inc = Run2{p, q};
            }
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                Run2 up{__shfl_up_sync(0xffffffffu, inc.p, d), __shfl_up_sync(0xffffffffu, inc.q, d)};
                if (lane >= d) inc = run2_combine(up, inc, double(kScanItems * d));
            }
            const Run2 tot{__shfl_sync(0xffffffffu, inc.p, 31), __shfl_sync(0xffffffffu, inc.q, 31)};
            Run2 excl{__shfl_up_sync(0xffffffffu, inc.p, 1), __shfl_up_sync(0xffffffffu, inc.q, 1)};
            if (lane == 0) excl = Run2{0.0, 0.0};
            const Run2 et = run2_combine(wpre, excl, double(kScanItems * lane));
            const Run2 full = run2_combine(tile_excl, et, double(k0));
            wpre = run2_combine(wpre, tot, double(32 * kScanItems));
            // Warp-uniform: the whole pass (256 samples) lies inside the grid and holds
            // neither P[n-1] nor P[n-2] (those go to the scalar block, slow path below).
            const uint64_t pass0 = tile * uint64_t(kScanTile) + uint64_t(j) * (32 * kScanItems);
            const bool pass_fast = uint32_t(j + 1) * (32 * kScanItems) <= m && pass0 + 32 * kScanItems + 1 < P.n;
            if (pass_fast) {
                // fast path: a full run of 8 samples, stored as 2 x 256 bits per table
                double p = 0.0, q = 0.0;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    double Ph[4], Qh[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int ii = half * 4 + i;
                        p += v[ii], q += p;
                        Ph[i] = full.p + p;
                        Qh[i] = fma(double(ii + 1), full.p, full.q) + q;
                    }
                    ptx::st_f64x4(st.P + idx0 + 4 * half, Ph[0], Ph[1], Ph[2], Ph[3]);
                    ptx::st_f64x4(st.Q + idx0 + 4 * half, Qh[0], Qh[1], Qh[2], Qh[3]);
                    if (replicate) {   // stage for the peer stores below
                        double* sp = stage + lane * kScanItems + 4 * half;
                        *reinterpret_cast<double4*>(sp) = make_double4(Ph[0], Ph[1], Ph[2], Ph[3]);
                        *reinterpret_cast<double4*>(sp + 32 * kScanItems) = make_double4(Qh[0], Qh[1], Qh[2], Qh[3]);
                    }
                }
                if (replicate) {
                    // Fused all-gather: the pass goes to every peer's tables in stores that
                    // are contiguous across the warp (1 KB per instruction).  Stores over
                    // NVLink are not merged by an L2 on the way out, so the lane-strided
                    // layout used for local HBM would leave as 32-byte packets.
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        const int o = i * 128 + lane * 4;
                        const double4 pv = *reinterpret_cast<const double4*>(stage + o);
                        const double4 qv = *reinterpret_cast<const double4*>(stage + 32 * kScanItems + o);
                        for (int r = 0; r < P.world; ++r) {
                            if (r == P.rank) continue;
                            ptx::st_f64x4(st.peerP[r] + pass0 + o, pv.x, pv.y, pv.z, pv.w);
                            ptx::st_f64x4(st.peerQ[r] + pass0 + o, qv.x, qv.y, qv.z, qv.w);
                        }
                    }
                    __syncwarp();
                }
                if (st.samples) {
                    ptx::st_f64x4(st.samples + idx0, v[0], v[1], v[2], v[3]);
                    ptx::st_f64x4(st.samples + idx0 + 4, v[4], v[5], v[6], v[7]);
                }
            } else {
                // ragged end of the grid, and the run that holds the final scalars
                double p = 0.0, q = 0.0;
#pragma unroll
                for (int i = 0; i < kScanItems; ++i) {
                    p += v[i], q += p;
                    const double Pi = full.p + p;
                    const double Qi = fma(double(i + 1), full.p, full.q) + q;
                    const uint64_t g = idx0 + i;
                    if (k0 + i < m) {
                        st.P[g] = Pi;
                        st.Q[g] = Qi;
                        if (st.samples) st.samples[g] = v[i];
                        if (replicate)
                            for (int r = 0; r < P.world; ++r)
                                if (r != P.rank) st.peerP[r][g] = Pi, st.peerQ[r][g] = Qi;
                    }
                    // The three scalars the callers print (4main.c:241 uses P[n-2]).
                    if (g + 1 == P.n || g + 2 == P.n) {
                        for (int r = 0; r < P.world; ++r) {
                            double* fin = st.peers[r]->scan_final;
                            if (g + 1 == P.n) {
                                ptx::st_relaxed_sys_f64(fin + 0, Pi);
                                ptx::st_relaxed_sys_f64(fin + 1, Qi);
                                if (P.n == 1) ptx::st_relaxed_sys_f64(fin + 2, 0.0);
                            } else {
                                ptx::st_relaxed_sys_f64(fin + 2, Pi);
                            }
                        }
                    }
                }
            }
        }
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
    // device -> host leg: the result block goes straight into pinned mapped host memory
    if (st.result_host) {
        __syncwarp();
        static_assert(sizeof(bq_scan_result) % 8 == 0 && sizeof(bq_scan_result) / 8 <= 32,
                      "scan result block = one word per lane");
        const unsigned long long* src = reinterpret_cast<const unsigned long long*>(st.result);
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(st.result_host);
        if (lane < int(sizeof(bq_scan_result) / 8)) dst[lane] = src[lane];
    }
}

template <class F, bool REPL>
cudaError_t launch_scan_fr(const ScanPlan& P, const ScanState& st, unsigned grid, cudaStream_t stream) {
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(grid, 1, 1);
    lc.blockDim = dim3(kScanThreads, 1, 1);
    lc.stream = stream;
    // staging for the coalesced peer stores of the fused all-gather
    lc.dynamicSmemBytes = REPL ? kScanStageBytes : 0;
    if (REPL) {
        // static table + staging exceeds the default 48 KB for the table integrand: opt in once per device
        static std::atomic<bool> done[64];   // zero-initialised; set once per device, any thread
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && !done[dev].load(std::memory_order_acquire)) {
            cudaError_t e = cudaFuncSetAttribute(scan_kernel<F, REPL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 int(kScanStageBytes));
            if (e != cudaSuccess) return e;
            done[dev].store(true, std::memory_order_release);
        }
    }
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    lc.attrs = attr;
    lc.numAttrs = 1;
    ScanParams prm;
    prm.P = P;
    prm.st = st;
    return cudaLaunchKernelEx(&lc, scan_kernel<F, REPL>, prm);
}

template <class F>
cudaError_t launch_scan_f(const ScanPlan& P, const ScanState& st, unsigned grid, cudaStream_t stream) {
    const bool replicate = P.world > 1 && (P.flags & kScanReplicate);
    return replicate ? launch_scan_fr<F, true>(P, st, grid, stream) : launch_scan_fr<F, false>(P, st, grid, stream);
}

// Co-resident CTAs per SM of both instantiations (the launch is cooperative).  The
// replicate kernel's static + staging shared memory exceeds the default 48 KB for the
// table integrand, so the opt-in has to precede the query or it reports zero.
template <class F>
int occupancy_f() {
    int plain = 0, repl = 0;
    cudaFuncSetAttribute(scan_kernel<F, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kScanStageBytes));
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&plain, scan_kernel<F, false>, kScanThreads, 0) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&repl, scan_kernel<F, true>, kScanThreads, kScanStageBytes) !=
            cudaSuccess) {
        cudaGetLastError();
        return 1;
    }
    const int n = plain < repl ? plain : repl;
    return n < 1 ? 1 : n;
}

}  // namespace

int scan_max_grid(int /*device*/, int sm_count) {
    int occ = occupancy_f<ProfileF>();
    const int o2 = occupancy_f<DisF>();
    if (o2 < occ) occ = o2;
    if (occ > 2) occ = 2;
    return occ * sm_count;
}

cudaError_t launch_scan(const ScanPlan& P, const ScanState& st, unsigned grid, cudaStream_t stream) {
    switch (P.integrand) {
        case BQ_F_SIN: return launch_scan_f<SinF>(P, st, grid, stream);
        case BQ_F_ACC: return launch_scan_f<AccF>(P, st, grid, stream);
        case BQ_F_VEL: return launch_scan_f<VelF>(P, st, grid, stream);
        case BQ_F_DIS: return launch_scan_f<DisF>(P, st, grid, stream);
        case BQ_F_PROFILE: return launch_scan_f<ProfileF>(P, st, grid, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace bq
