// The fused cumulative-integration kernel: ONE launch per GPU per call, no grid-wide barrier.
//
// Reference capability (4main.c:76-225): fill InterpProfile, per-rank running
// sum, gather to root, serial carry fix-up on the root, MPI_Bcast of 144 MB —
// twice.  Here every WARP of a persistent grid owns whole tiles (2048 samples) and does, per tile,
//
//   A   evaluate f on the fly (no sample array is ever materialised) and reduce the tile to its
//       aggregate (sum v, sum of local running sums); publish it as a descriptor
//   D   "duties" of the last tile of a group / supergroup / segment: fold the aggregates of its
//       group into the group total, the group totals into the supergroup prefix, and — for the
//       last tile of a segment — store the segment total into every rank's comm block over
//       NVLink                                            (replaces 4main.c:134-155, 200-219)
//   B   read the three descriptors its exclusive prefix is made of (scan_plan.h: its prefix inside
//       the group, the group's prefix inside the segment, the segment base), re-evaluate
//       the tile, scan it and stream P and Q out in 256-bit stores — into every peer's table
//       as well when replicate is on (the fused all-gather that replaces MPI_Bcast,
//       4main.c:157,221)
//
// A descriptor is a few "LL" words — 32 bits of payload plus the 32-bit generation of the launch
// per 8-byte word — so publishing is a plain store and waiting is polling the words themselves:
// no fence, no flag, no counter, nothing to clear between calls.  A and D only ever wait for A / D
// of EARLIER tiles; B waits for A / D of tiles at most 1023 further on, which every warp has been
// through before it starts that B (the A parts run ahead, see the tile loop); all warps are
// resident (cooperative launch), so the waits end.
//
// One GPU runs a single pass: A and D one round of the grid ahead of B, so the evaluation and the
// prefix chain of the next round overlap the stores of this one.  With several GPUs a rank needs
// the lower ranks' segment totals before its first store, so A + D cover the whole slice first.
// The prefix tree is the same fixed shape either way, so the tables are bit-identical for every
// world size and match cpu_scan.cpp bit for bit.
//
// A tile is 8 passes of 32 lanes x 8 samples.  The arithmetic tree of a tile (scan_core.cuh) is
// that of a 256-thread block scan; one warp walks it pass by pass.
#include <cuda_runtime.h>
#include <stdlib.h>

#include <atomic>
#include <type_traits>

#include "ptx.cuh"
#include "quad_core.cuh"
#include "scan_core.cuh"
#include "scan_launch.h"

#ifndef BQ_SCAN_EXTRA_AHEAD
#define BQ_SCAN_EXTRA_AHEAD 1   // grid rounds the A parts run ahead beyond the 1024 tiles a B part may depend on
#endif
#ifndef BQ_SCAN_POLL_NS
#define BQ_SCAN_POLL_NS 100
#endif
#ifndef BQ_SCAN_B_PASSES
#define BQ_SCAN_B_PASSES 1      // passes of a tile whose lane scans the B part issues interleaved (1, 2 or 4); 2 spills at 128 registers and is 12 % slower
#endif
#ifndef BQ_SCAN_MIN_CTAS
#define BQ_SCAN_MIN_CTAS 2      // resident CTAs per SM the register budget is set for (2 -> 128 registers)
#endif

namespace bq {

namespace {

constexpr int kPasses = kScanThreads / 32;          // passes of a tile: part of the arithmetic tree (scan_plan.h)
// Threads of a CTA of the launch — independent of the tile tree (a warp owns a tile whatever CTA it sits in).
#ifndef BQ_SCAN_CTA_THREADS
#define BQ_SCAN_CTA_THREADS 256
#endif
constexpr int kCtaThreads = BQ_SCAN_CTA_THREADS;
constexpr int kBPasses = BQ_SCAN_B_PASSES;
static_assert(kBPasses == 1 || kBPasses == 2 || kBPasses == 4, "BQ_SCAN_B_PASSES divides the 8 passes of a tile");
constexpr int kWarps = kCtaThreads / 32;
constexpr size_t kScanStageBytes = size_t(kWarps) * 2 * 32 * kScanItems * sizeof(double);   // 32 KB at 8 warps

struct ScanSmem {
    uint64_t gen;
    int is_last;
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
__device__ __forceinline__ Run shfl_run(const Run& r, int src) {
    Run o;
    o.p.hi = __shfl_sync(0xffffffffu, r.p.hi, src);
    o.p.lo = __shfl_sync(0xffffffffu, r.p.lo, src);
    o.q.hi = __shfl_sync(0xffffffffu, r.q.hi, src);
    o.q.lo = __shfl_sync(0xffffffffu, r.q.lo, src);
    o.cnt = __shfl_sync(0xffffffffu, r.cnt, src);
    return o;
}
// Kogge-Stone inclusive scan over the lanes (lanes_inclusive() of cpu_scan.cpp).
__device__ __forceinline__ Run lanes_inclusive(Run inc) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const Run up = shfl_up_run(inc, d);
        if (lane >= d) inc = run_combine(up, inc);
    }
    return inc;
}

// A (dd P, dd Q) pair as 8 LL words.
template <bool SYS>
__device__ __forceinline__ void run_put(uint64_t* dst, const Run& r, uint32_t flag) {
    ptx::ll_put<SYS>(dst + 0, r.p.hi, flag);
    ptx::ll_put<SYS>(dst + 2, r.p.lo, flag);
    ptx::ll_put<SYS>(dst + 4, r.q.hi, flag);
    ptx::ll_put<SYS>(dst + 6, r.q.lo, flag);
}
template <bool SYS>
__device__ __forceinline__ bool run_get(const uint64_t* src, uint32_t flag, Run& r) {
    double a, b, c, d;
    const bool ok = ptx::ll_get<SYS>(src + 0, flag, a) & ptx::ll_get<SYS>(src + 2, flag, b) &
                    ptx::ll_get<SYS>(src + 4, flag, c) & ptx::ll_get<SYS>(src + 6, flag, d);
    if (ok) r.p = dd{a, b}, r.q = dd{c, d};
    return ok;
}

// Warp-collective end-of-wait test for the polling loops: done when every lane has what it needs,
// or when this warp (or any other: abort word) has waited longer than the time-out.
struct Waiter {
    const uint64_t* abort_word;
    uint64_t* abort_word_w;
    uint64_t gen, limit_ns, t0;
    uint32_t spins;
    bool failed;
    __device__ __forceinline__ Waiter(uint64_t* aw, uint64_t g, uint64_t limit)
        : abort_word(aw), abort_word_w(aw), gen(g), limit_ns(limit), t0(0), spins(0), failed(false) {}
    __device__ __forceinline__ bool done(bool ok) {
        if (__all_sync(0xffffffffu, ok)) return true;
        __nanosleep(BQ_SCAN_POLL_NS);      // a failed poll yields its issue slots to the warps that have work
        if ((++spins & 31u) == 0u) {
            if (!t0) t0 = ptx::globaltimer();
            const bool late = ptx::globaltimer() - t0 > limit_ns;
            const bool aborted = ptx::ld_acquire_gpu(abort_word) == gen;
            if (__any_sync(0xffffffffu, late || aborted)) {
                if ((threadIdx.x & 31) == 0) ptx::st_release_gpu(abort_word_w, gen);
                failed = true;
                return true;
            }
        }
        return false;
    }
    __device__ __forceinline__ void rearm() { t0 = 0, spins = 0; }
};

struct ScanParams {
    ScanPlan P;        // rides in the kernel parameter buffer
    ScanState st;
};

// Where tile `tile` sits in the prefix tree of its segment.
struct TilePos {
    int seg;
    uint64_t j, T;          // index inside the segment, tiles of the segment
    int k, gi;              // index inside the group, of the group inside its supergroup
    uint64_t g, sg;         // group / supergroup index inside the segment
    bool last_in_seg;
};
__device__ __forceinline__ TilePos tile_pos(const ScanPlan& P, uint64_t tile) {
    TilePos t;
    t.seg = scan_segment_of(P, tile);
    t.j = tile - P.seg_tile[t.seg];
    t.T = P.seg_tile[t.seg + 1] - P.seg_tile[t.seg];
    t.g = t.j / kScanGroup;
    t.k = int(t.j % kScanGroup);
    t.sg = t.g / kScanSuper;
    t.gi = int(t.g % kScanSuper);
    t.last_in_seg = t.j + 1 == t.T;
    return t;
}

// Two resident CTAs per SM (128 registers): measured best on B200 — the register room lets
// the compiler overlap the passes of a tile; 3 and 4 CTAs/SM (80 / 64 registers) both lost.
// REPL = the fused all-gather (replicate) is compiled in; a separate instantiation so the
// single-GPU / slice-only kernel carries none of its registers or staging.
template <class F, bool REPL>
__global__ void __launch_bounds__(kCtaThreads, BQ_SCAN_MIN_CTAS) scan_kernel(const __grid_constant__ ScanParams prm) {
    const ScanPlan& P = prm.P;
    const ScanState& st = prm.st;
    __shared__ ScanSmem sm;
    constexpr bool kIsProfile = std::is_same<F, ProfileF>::value;
    __shared__ __align__(16) double sTable[kIsProfile ? kMaxTable : 2];
    __shared__ __align__(8) uint64_t sBar;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (blockIdx.x == 0 && tid == 0) st.result->t_start_ns = ptx::globaltimer();
    if (tid == 0) {
        // generation of this launch: one more than the number of scans this context has completed
        // (written by the last CTA of the previous launch, ordered by the stream)
        sm.gen = st.comm->scan_epoch + 1;
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
    const uint32_t flag = uint32_t(gen % 0xffffffffull) + 1u;     // never 0
    const int par = int(gen & 1);
    const bool multi = P.world > 1;
    const uint64_t gwarp = uint64_t(blockIdx.x) * kWarps + warp;
    const uint64_t nwarps = uint64_t(gridDim.x) * kWarps;
    const bool replicate = REPL && multi && (P.flags & kScanReplicate);
    extern __shared__ __align__(32) double sStage[];   // replicate only: [warp][P 256 | Q 256]
    double* const stage = sStage + warp * (2 * 32 * kScanItems);
    // (a Waiter is built where it is used: nothing of it stays live across the tile loop; a time-out
    // is remembered in the abort word, not in a register)

    // Totals of owned segments that hold no tile (tiny n): nobody's duty, published up front.
    if (gwarp == 0 && lane < P.world && !(P.flags & kScanSkipPublish)) {
        for (uint32_t s = P.seg_lo; s < P.seg_hi; ++s)
            if (P.seg_tile[s + 1] == P.seg_tile[s])
                run_put<true>(&st.peers[lane]->scan_ll[par][s][0], run_identity(), flag);
    }

    // ---- A + duties of one tile ------------------------------------------------------------
    auto do_a = [&](uint64_t tile) {
        Waiter wait(st.abort_word, gen, limit);
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
        uint64_t* mine = st.a_ll + 4 * (tile - P.tile_lo);
        if (lane == 0) {
            ptx::ll_put<false>(mine + 0, S, flag);
            ptx::ll_put<false>(mine + 2, W, flag);
        }
        const TilePos tp = tile_pos(P, tile);
        if (tp.k != kScanGroup - 1 && !tp.last_in_seg) return;

        // -- group duty: inclusive scan over the group's aggregates; every later tile of the group
        //    gets its in-group prefix Ascan[k-1] as a descriptor, the group its total GT(g)
        Run r = run_identity();
        {
            double s = 0.0, w = 0.0;
            bool ok = lane >= tp.k;
            const uint64_t* src = mine - 4 * (tp.k - lane);     // tile - k + lane
            wait.rearm();
            for (;;) {
                if (!ok) ok = ptx::ll_get<false>(src, flag, s) && ptx::ll_get<false>(src + 2, flag, w);
                if (wait.done(ok)) break;
            }
            if (lane < tp.k) r = run_from_tile(s, w, double(kScanTile));     // predecessors are full tiles
            if (lane == tp.k) r = run_from_tile(S, W, double(m));
        }
        const Run ainc = lanes_inclusive(r);
        if (lane < tp.k) run_put<false>(st.aex_ll + 8 * (tile - P.tile_lo - tp.k + lane + 1), ainc, flag);
        const Run gt = shfl_run(ainc, tp.k);
        const uint64_t gidx = P.seg_group0[tp.seg] - P.seg_group0[P.seg_lo] + tp.g;
        if (lane == 0) run_put<false>(st.gt_ll + 8 * gidx, gt, flag);
        if (tp.gi != kScanSuper - 1 && !tp.last_in_seg) return;

        // -- supergroup duty: scan over its group totals; every group gets its in-segment prefix
        //    head(g) = SGP(sg-1) (+) GTscan[gi-1], the supergroup its closing prefix SGP(sg)
        Run q = run_identity();
        {
            Run got = run_identity();
            bool ok = lane >= tp.gi;
            const uint64_t* src = st.gt_ll + 8 * (gidx - tp.gi + lane);
            wait.rearm();
            for (;;) {
                if (!ok) ok = run_get<false>(src, flag, got);
                if (wait.done(ok)) break;
            }
            if (lane < tp.gi) q = got, q.cnt = double(kScanGroup) * double(kScanTile);   // predecessors are full groups
            if (lane == tp.gi) q = gt;
        }
        const Run ginc = lanes_inclusive(q);
        const uint64_t sgidx = P.seg_super0[tp.seg] - P.seg_super0[P.seg_lo] + tp.sg;
        Run prev = run_identity();
        if (tp.sg > 0) {
            bool ok = false;
            wait.rearm();
            for (;;) {
                if (!ok) ok = run_get<false>(st.sgp_ll + 8 * (sgidx - 1), flag, prev);
                if (wait.done(ok)) break;
            }
            prev.cnt = double(tp.sg) * double(kScanSuper) * double(kScanGroup) * double(kScanTile);
        }
        {
            Run gexcl = shfl_up_run(ginc, 1);
            if (lane == 0) gexcl = run_identity();
            const Run head = run_combine(prev, gexcl);
            if (lane <= tp.gi) run_put<false>(st.head_ll + 8 * (gidx - tp.gi + lane), head, flag);
        }
        const Run sgp = run_combine(prev, shfl_run(ginc, tp.gi));
        if (lane == 0) run_put<false>(st.sgp_ll + 8 * sgidx, sgp, flag);
        if (!tp.last_in_seg) return;

        // -- segment duty: the segment total goes to every rank's comm block (NVLink stores)
        if (lane < P.world && !(P.flags & kScanSkipPublish))
            run_put<true>(&st.peers[lane]->scan_ll[par][tp.seg][0], sgp, flag);
    };

    // ---- B of one tile: prefix from descriptors, scan, store ----------------------------------
    auto do_b = [&](uint64_t tile) {
        Waiter wait(st.abort_word, gen, limit);
        const uint32_t m = scan_tile_count(P, tile);
        const TilePos tp = tile_pos(P, tile);
        const uint64_t gidx = P.seg_group0[tp.seg] - P.seg_group0[P.seg_lo] + tp.g;
        Run base = run_identity();
        if (tp.j == 0) {
            // first tile of a segment: segbase(s) = in-order fold of the totals of segments 0..s-1,
            // which every rank stored into this comm block; published for the segment's other tiles
            Run mine = run_identity();
            bool ok = lane >= tp.seg;
            wait.rearm();
            for (;;) {
                if (!ok) ok = run_get<true>(&st.comm->scan_ll[par][lane & (kSegments - 1)][0], flag, mine);
                if (wait.done(ok)) break;
            }
            const int ls = lane & (kSegments - 1);
            mine.cnt = scan_samples_in_tiles(P, P.seg_tile[ls], P.seg_tile[ls + 1]);
            for (int s = 0; s < tp.seg; ++s) base = run_combine(base, shfl_run(mine, s));
            if (lane == 0) run_put<false>(st.segbase_ll + 8 * tp.seg, base, flag);
        }
        // exclusive prefix of the tile: segbase (+) [head(g) (+) Ascan[k-1]] — three descriptors, written
        // by the first tile of the segment and by the last tiles of the supergroup and of the group
        Run2 tile_excl;
        {
            Run aexcl = run_identity(), head = run_identity();
            bool ok_a = tp.k == 0, ok_h = false, ok_b = tp.j == 0;
            const uint64_t* asrc = st.aex_ll + 8 * (tile - P.tile_lo);
            const uint64_t* hsrc = st.head_ll + 8 * gidx;
            wait.rearm();
            for (;;) {
                if (!ok_a) ok_a = run_get<false>(asrc, flag, aexcl);
                if (!ok_h) ok_h = run_get<false>(hsrc, flag, head);
                if (!ok_b) ok_b = run_get<false>(st.segbase_ll + 8 * tp.seg, flag, base);
                if (wait.done(ok_a && ok_h && ok_b)) break;
            }
            aexcl.cnt = double(tp.k) * double(kScanTile);
            head.cnt = double(tp.g) * double(kScanGroup) * double(kScanTile);
            const Run pre = run_combine(head, aexcl);
            tile_excl = scan_tile_exclusive(base, pre);
        }
        // One pass: the values of this lane's run and the run's local totals ...
        auto pass_values = [&](int j, double (&v)[kScanItems]) -> Run2 {
            const uint32_t k0 = uint32_t(j * 32 + lane) * kScanItems;
            scan_thread_values(f, P, tile * uint64_t(kScanTile) + k0, k0, m, v);
            double p = 0.0, q = 0.0;
#pragma unroll
            for (int i = 0; i < kScanItems; ++i) p += v[i], q += p;
            return Run2{p, q};
        };
        // ... and, once the run's exclusive prefix `full` is known, its 8 table entries.
        auto emit = [&](int j, const double (&v)[kScanItems], const Run2& full) {
            const uint32_t k0 = uint32_t(j * 32 + lane) * kScanItems;
            const uint64_t idx0 = tile * uint64_t(kScanTile) + k0;
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
                        if (st.mcP) {
                            // NVLS: one store per 16 bytes to the multicast address; the switch
                            // writes every rank's copy (this one's too — harmless, same values)
                            ptx::multimem_st_f64x2(st.mcP + pass0 + o, pv.x, pv.y);
                            ptx::multimem_st_f64x2(st.mcP + pass0 + o + 2, pv.z, pv.w);
                            ptx::multimem_st_f64x2(st.mcQ + pass0 + o, qv.x, qv.y);
                            ptx::multimem_st_f64x2(st.mcQ + pass0 + o + 2, qv.z, qv.w);
                        } else {
                            for (int r = 0; r < P.world; ++r) {
                                if (r == P.rank) continue;
                                ptx::st_f64x4(st.peerP[r] + pass0 + o, pv.x, pv.y, pv.z, pv.w);
                                ptx::st_f64x4(st.peerQ[r] + pass0 + o, qv.x, qv.y, qv.z, qv.w);
                            }
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
                            uint64_t* fin = &st.peers[r]->scan_final_ll[par][0];
                            if (g + 1 == P.n) {
                                ptx::ll_put<true>(fin + 0, Pi, flag);
                                ptx::ll_put<true>(fin + 2, Qi, flag);
                                if (P.n == 1) ptx::ll_put<true>(fin + 4, 0.0, flag);
                            } else {
                                ptx::ll_put<true>(fin + 4, Pi, flag);
                            }
                        }
                    }
                }
            }
        };
        Run2 wpre{0.0, 0.0};   // fold of the passes before this one ("warps before mine")
        // Two passes at a time: their lane scans (five dependent shuffle steps each) are independent
        // and are issued interleaved; the arithmetic is that of the pass-by-pass walk.
#pragma unroll 1
        for (int j = 0; j < kPasses; j += kBPasses) {
            double v[kBPasses][kScanItems];
            Run2 inc[kBPasses];
#pragma unroll
            for (int u = 0; u < kBPasses; ++u) inc[u] = pass_values(j + u, v[u]);
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                Run2 up[kBPasses];
#pragma unroll
                for (int u = 0; u < kBPasses; ++u)
                    up[u] = Run2{__shfl_up_sync(0xffffffffu, inc[u].p, d), __shfl_up_sync(0xffffffffu, inc[u].q, d)};
#pragma unroll
                for (int u = 0; u < kBPasses; ++u)
                    if (lane >= d) inc[u] = run2_combine(up[u], inc[u], double(kScanItems * d));
            }
            Run2 tot[kBPasses], excl[kBPasses];
#pragma unroll
            for (int u = 0; u < kBPasses; ++u) {
                tot[u] = Run2{__shfl_sync(0xffffffffu, inc[u].p, 31), __shfl_sync(0xffffffffu, inc[u].q, 31)};
                excl[u] = Run2{__shfl_up_sync(0xffffffffu, inc[u].p, 1), __shfl_up_sync(0xffffffffu, inc[u].q, 1)};
                if (lane == 0) excl[u] = Run2{0.0, 0.0};
            }
#pragma unroll
            for (int u = 0; u < kBPasses; ++u) {
                const uint32_t k0 = uint32_t((j + u) * 32 + lane) * kScanItems;
                const Run2 et = run2_combine(wpre, excl[u], double(kScanItems * lane));
                const Run2 full = run2_combine(tile_excl, et, double(k0));
                wpre = run2_combine(wpre, tot[u], double(32 * kScanItems));
                emit(j + u, v[u], full);
            }
        }
    };

    // ---- the tile loop -----------------------------------------------------------------------
    // A tile's B part reads descriptors that the LAST tiles of its group and supergroup write, up
    // to 1023 tiles further on, so the A parts (and duties) run ahead of the B parts: by `ahead`
    // rounds of the grid on one GPU (>= 1024 tiles, so every descriptor a B part waits for belongs
    // to an A part that its warp has already been through — the waits always end — and the
    // evaluation of the next round overlaps the stores of this one); by the whole slice with
    // several GPUs, whose first store also needs the lower ranks' segment totals.
    {
        const uint64_t first = P.tile_lo + gwarp;
        // (+1: a whole round of slack, so that what a B part reads was written an iteration ago and
        // not by the A parts running next to it — measured: 17 % of the stall samples sat in those waits)
        uint64_t ahead = (uint64_t(kScanGroup) * kScanSuper + nwarps - 1) / nwarps + BQ_SCAN_EXTRA_AHEAD;
        if (multi) ahead = (P.tile_hi - P.tile_lo + nwarps - 1) / nwarps + 1;
#pragma unroll 1
        for (uint64_t it = 0;; ++it) {
            const uint64_t a_tile = first + it * nwarps;
            if (a_tile < P.tile_hi) do_a(a_tile);
            if (it < ahead) continue;
            const uint64_t b_tile = first + (it - ahead) * nwarps;
            if (b_tile >= P.tile_hi) break;
            do_b(b_tile);
        }
    }

    // ---- end: the last CTA collects the final scalars and writes the result ------------------
    __syncthreads();
    if (tid == 0) {
        if (multi) ptx::fence_sys(); else __threadfence();
        sm.is_last = (atomicAdd(st.ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!sm.is_last || warp != 0) return;
    if (multi) ptx::fence_sys(); else __threadfence();
    Waiter wait(st.abort_word, gen, limit);
    if (replicate) {
        // Fused all-gather: every rank tells every rank "my stores into your tables are complete"
        // (a release store after the system fence above) and waits for the same word from everyone;
        // when the kernel retires, this GPU's tables are whole.
        if (lane < P.world && !(P.flags & kScanSkipPublish))
            ptx::st_release_sys_u64(&st.peers[lane]->scan_done_ll[par][P.rank], uint64_t(flag));
        bool ok = lane >= P.world;
        wait.rearm();
        for (;;) {
            if (!ok) ok = ptx::ld_acquire_sys(&st.comm->scan_done_ll[par][lane & (BQ_MAX_WORLD - 1)]) == uint64_t(flag);
            if (wait.done(ok)) break;
        }
    }
    // P[n-1], Q[n-1], P[n-2]: stored by whichever rank holds the end of the grid into every comm block
    double fin = 0.0;
    {
        bool ok = lane >= 3;
        wait.rearm();
        for (;;) {
            if (!ok) ok = ptx::ll_get<true>(&st.comm->scan_final_ll[par][2 * (lane & 3)], flag, fin);
            if (wait.done(ok)) break;
        }
    }
    const double p_last = __shfl_sync(0xffffffffu, fin, 0), q_last = __shfl_sync(0xffffffffu, fin, 1),
                 p_second = __shfl_sync(0xffffffffu, fin, 2);
    if (lane == 0) {
        const bool aborted = ptx::ld_acquire_gpu(st.abort_word) == gen;
        bq_scan_result* R = st.result;
        R->p_last = p_last;
        R->q_last = q_last;
        R->p_second_last = p_second;
        R->slice_begin = P.tile_lo * uint64_t(kScanTile);
        uint64_t e = P.tile_hi * uint64_t(kScanTile);
        R->slice_end = e < P.n ? e : P.n;
        R->status = (aborted || wait.failed) ? BQ_ERR_TIMEOUT : BQ_OK;
        R->world = P.world;
        R->rank = P.rank;
        R->reserved[0] = R->reserved[1] = 0;
        *st.ticket = 0u;
        st.comm->scan_epoch = gen;     // generation completed: the next launch uses gen + 1
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
    lc.blockDim = dim3(kCtaThreads, 1, 1);
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
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&plain, scan_kernel<F, false>, kCtaThreads, 0) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&repl, scan_kernel<F, true>, kCtaThreads, kScanStageBytes) !=
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
    if (occ > BQ_SCAN_MIN_CTAS) occ = BQ_SCAN_MIN_CTAS;
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
