// Host side of the fused scan: plan, scratch management, launches.
#include <cuda_runtime.h>
#include <string.h>

#include <vector>

#include "ctx.h"
#include "host_common.h"
#include "nvtx_range.h"
#include "scan_launch.h"

namespace bq {

#define BQ_CUDA(expr)                                                                      \
    do {                                                                                   \
        cudaError_t e_ = (expr);                                                           \
        if (e_ != cudaSuccess) {                                                           \
            set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__,    \
                      __LINE__);                                                           \
            return BQ_ERR_CUDA;                                                            \
        }                                                                                  \
    } while (0)

namespace {
struct DevGuard {
    int prev = -1;
    explicit DevGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
    }
    ~DevGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};
}  // namespace

constexpr size_t kScanMiscWords = 128;   // 64 segment-base words, abort word, spare

int scan_ctx_init(bq_ctx* c) {
    BQ_CUDA(cudaMalloc(&c->d_scan_misc, sizeof(uint64_t) * kScanMiscWords));
    BQ_CUDA(cudaMemset(c->d_scan_misc, 0, sizeof(uint64_t) * kScanMiscWords));
    c->scan_grid = scan_max_grid(c->device, c->sm_count);
    return BQ_OK;
}

void scan_ctx_free(bq_ctx* c) {
    cudaFree(c->d_scan_misc);
    cudaFree(c->d_scan_scratch);
}

// Descriptor words are valid when they carry the running launch's generation, and generations
// restart when the comm block is reset: the scratch goes back to zero with it (0 is never valid).
int scan_clear_scratch(bq_ctx* c) {
    if (c->d_scan_misc) BQ_CUDA(cudaMemset(c->d_scan_misc, 0, sizeof(uint64_t) * kScanMiscWords));
    if (c->d_scan_scratch) BQ_CUDA(cudaMemset(c->d_scan_scratch, 0, scan_scratch_layout(c->scan_tiles_cap).total));
    return BQ_OK;
}

static int scan_reserve(bq_ctx* c, uint64_t tiles, cudaStream_t stream) {
    if (tiles <= c->scan_tiles_cap) return BQ_OK;
    // Growing happens outside any timed path (first call of a given size) and never inside a
    // stream capture: cudaMalloc / cudaFree are not capturable.  bq_ctx_reserve_scan() sizes the
    // scratch ahead of a capture.
    cudaStreamCaptureStatus cap_status = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &cap_status) == cudaSuccess && cap_status != cudaStreamCaptureStatusNone) {
        set_error("cumulative: the scan scratch must grow for this n, which cannot happen inside a stream "
                  "capture; call bq_ctx_reserve_scan(ctx, n) before capturing");
        return BQ_ERR_INVALID;
    }
    BQ_CUDA(cudaStreamSynchronize(c->last_stream ? c->last_stream : c->stream));
    cudaFree(c->d_scan_scratch);
    c->d_scan_scratch = nullptr;
    c->scan_tiles_cap = 0;
    uint64_t cap = 16384;
    while (cap < tiles) cap *= 2;
    const size_t bytes = scan_scratch_layout(cap).total;
    BQ_CUDA(cudaMalloc(&c->d_scan_scratch, bytes));
    BQ_CUDA(cudaMemset(c->d_scan_scratch, 0, bytes));
    c->scan_tiles_cap = cap;
    return BQ_OK;
}

static int scan_enqueue(bq_ctx* c, const bq_scan_job* job, int rank, int world, double* dP,
                        double* dQ, double* dS, cudaStream_t stream) {
    if (!dP || !dQ) {
        set_error("cumulative: P and Q output buffers are required");
        return BQ_ERR_INVALID;
    }
    if ((uintptr_t(dP) | uintptr_t(dQ) | uintptr_t(dS)) & 31) {
        set_error("cumulative: output buffers must be 32-byte aligned");
        return BQ_ERR_INVALID;
    }
    NvtxRange range("b200quad::cumulative");
    // The plan travels in the kernel parameter buffer: no staging copy.
    ScanPlan plan;
    ScanPlan* hp = &plan;
    int rc = make_scan_plan(*job, rank, world, c->table_len, hp);
    if (rc != BQ_OK) return rc;
    hp->timeout_ms = c->timeout_ms;
    if (c->fault_flags) hp->flags |= kScanSkipPublish;
    c->fault_flags = 0;
    const bool replicate = world > 1 && (hp->flags & kScanReplicate);
    if (world > 1) {
        for (int r = 0; r < world; ++r) {
            if (r != rank && !c->peers[r]) {
                set_error("rank %d: peer %d is not attached (world=%d)", rank, r, world);
                return BQ_ERR_INVALID;
            }
            if (replicate && r != rank && (!c->scan_peerP[r] || !c->scan_peerQ[r])) {
                set_error("replicate: peer tables of rank %d not set (bq_ctx_set_scan_peers)", r);
                return BQ_ERR_INVALID;
            }
        }
    }
    rc = scan_reserve(c, hp->tile_hi - hp->tile_lo, stream);
    if (rc != BQ_OK) return rc;

    ScanState st;
    memset(&st, 0, sizeof(st));
    const ScanScratchLayout L = scan_scratch_layout(c->scan_tiles_cap);
    st.a_ll = reinterpret_cast<uint64_t*>(c->d_scan_scratch + L.a_off);
    st.aex_ll = reinterpret_cast<uint64_t*>(c->d_scan_scratch + L.aex_off);
    st.gt_ll = reinterpret_cast<uint64_t*>(c->d_scan_scratch + L.gt_off);
    st.head_ll = reinterpret_cast<uint64_t*>(c->d_scan_scratch + L.head_off);
    st.sgp_ll = reinterpret_cast<uint64_t*>(c->d_scan_scratch + L.sgp_off);
    st.segbase_ll = c->d_scan_misc;
    st.abort_word = c->d_scan_misc + 64;
    st.ticket = c->d_ticket + 16;
    st.result = c->d_scan_result;
    st.result_host = c->h_scan_result_dev;   // the kernel itself writes the result block to the host
    st.comm = c->d_comm;
    for (int r = 0; r < BQ_MAX_WORLD; ++r) st.peers[r] = c->peers[r];
    st.peers[rank] = c->d_comm;
    st.mc_comm = (c->use_nvls && world == c->world) ? c->mc_comm : nullptr;
    st.table = c->d_table;
    // Tables are indexed globally inside the kernel.  A replicating rank holds
    // the whole table; otherwise the buffer starts at this rank's slice.
    const uint64_t slice_begin = hp->tile_lo * uint64_t(kScanTile);
    const uint64_t shift = replicate ? 0 : slice_begin;
    st.P = dP - shift;
    st.Q = dQ - shift;
    st.samples = dS ? dS - shift : nullptr;
    for (int r = 0; r < BQ_MAX_WORLD; ++r) {
        st.peerP[r] = replicate ? c->scan_peerP[r] : nullptr;
        st.peerQ[r] = replicate ? c->scan_peerQ[r] : nullptr;
    }
    const bool mc_tables = replicate && c->use_nvls && c->scan_mcP && c->scan_mcQ;
    st.mcP = mc_tables ? c->scan_mcP : nullptr;
    st.mcQ = mc_tables ? c->scan_mcQ : nullptr;
    uint64_t tiles = hp->tile_hi - hp->tile_lo;
    unsigned grid = unsigned(c->scan_grid);
    if (tiles < grid) grid = unsigned(tiles ? tiles : 1);
    rc = order_after_previous(c, stream);
    if (rc != BQ_OK) return rc;
    BQ_CUDA(launch_scan(*hp, st, grid, stream));
    c->launches += 1;
    c->last_stream = stream;
    return BQ_OK;
}

}  // namespace bq

using namespace bq;

extern "C" {

int bq_ctx_set_scan_peers(bq_ctx* c, double* const* peerP, double* const* peerQ) {
    if (!c) return BQ_ERR_INVALID;
    for (int r = 0; r < BQ_MAX_WORLD; ++r) {
        c->scan_peerP[r] = peerP ? peerP[r] : nullptr;
        c->scan_peerQ[r] = peerQ ? peerQ[r] : nullptr;
    }
    return BQ_OK;
}

int bq_ctx_reserve_scan(bq_ctx* c, uint64_t n) {
    if (!c || n < 1 || n > kScanMaxN) return BQ_ERR_INVALID;
    DevGuard guard(c->device);
    return scan_reserve(c, (n + kScanTile - 1) / kScanTile, c->stream);
}

int bq_cumulative_async(bq_ctx* c, const bq_scan_job* job, double* dP, double* dQ, double* dS,
                        void* stream) {
    if (!c || !job) return BQ_ERR_INVALID;
    DevGuard guard(c->device);
    cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
    return scan_enqueue(c, job, c->rank, c->world, dP, dQ, dS, s);
}

int bq_scan_result_fetch(bq_ctx* c, bq_scan_result* out) {
    if (!c || !out) return BQ_ERR_INVALID;
    DevGuard guard(c->device);
    BQ_CUDA(cudaStreamSynchronize(c->last_stream ? c->last_stream : c->stream));
    *out = *c->h_scan_result;
    return out->status;
}

int bq_cumulative(bq_ctx* c, const bq_scan_job* job, double* dP, double* dQ, double* dS,
                  bq_scan_result* out) {
    if (!c || !job || !out) return BQ_ERR_INVALID;
    DevGuard guard(c->device);
    BQ_CUDA(cudaEventRecord(c->ev0, c->stream));
    int rc = scan_enqueue(c, job, c->rank, c->world, dP, dQ, dS, c->stream);
    if (rc != BQ_OK) return rc;
    BQ_CUDA(cudaEventRecord(c->ev1, c->stream));
    BQ_CUDA(cudaStreamSynchronize(c->stream));
    *out = *c->h_scan_result;
    float ms = 0.f;
    BQ_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    out->device_ms = ms;
    if (out->status != BQ_OK) set_error("scan kernel reported status %d", out->status);
    return out->status;
}

int bq_world_cumulative(bq_world* w, const bq_scan_job* job, double* const* dP, double* const* dQ,
                        bq_scan_result* out) {
    if (!w || !job || !dP || !dQ) return BQ_ERR_INVALID;
    const int n = bq_world_size(w);
    if (job->replicate && n > 1) {
        double* mcP = bq_world_table_multicast(w, dP[0]);
        double* mcQ = bq_world_table_multicast(w, dQ[0]);
        for (int r = 0; r < n; ++r) {
            int rc = bq_ctx_set_scan_peers(bq_world_ctx(w, r), dP, dQ);
            if (rc != BQ_OK) return rc;
            bq_world_ctx(w, r)->scan_mcP = mcP;     // null unless the tables came from bq_world_alloc_tables
            bq_world_ctx(w, r)->scan_mcQ = mcQ;
        }
    }
    for (int r = 0; r < n; ++r) {
        bq_ctx* c = bq_world_ctx(w, r);
        DevGuard guard(c->device);
        BQ_CUDA(cudaEventRecord(c->ev0, c->stream));
        int rc = scan_enqueue(c, job, r, n, dP[r], dQ[r], nullptr, c->stream);
        if (rc != BQ_OK) return rc;
        BQ_CUDA(cudaEventRecord(c->ev1, c->stream));
    }
    int worst = BQ_OK;
    for (int r = 0; r < n; ++r) {
        bq_ctx* c = bq_world_ctx(w, r);
        DevGuard guard(c->device);
        BQ_CUDA(cudaStreamSynchronize(c->stream));
        float ms = 0.f;
        BQ_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
        c->h_scan_result->device_ms = ms;
        if (out) out[r] = *c->h_scan_result;
        if (c->h_scan_result->status != BQ_OK) worst = c->h_scan_result->status;
    }
    return worst;
}

}  // extern "C"
