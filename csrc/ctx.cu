// CUDA host runtime: contexts, pinned result blocks, peer wiring, launches.
//
// The reference's host side is three main()s that cudaMalloc/cudaMemcpy 144 MB
// of uninitialised memory per run and never check an error
// (cintegrate.cu:111-133).  Here one context per device owns every buffer for
// its lifetime; a call is exactly one kernel launch: the plan rides down in
// the launch's parameter buffer and the kernel writes the result block into
// pinned, mapped host memory itself.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "ctx.h"
#include "drv.h"
#include "host_common.h"
#include "nvtx_range.h"
#include "quad_launch.h"
#include "scan_launch.h"
#include "symm.h"

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

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

}  // namespace bq

namespace bq {
static int ctx_init(bq_ctx* c, int device);
static int ctx_adopt_comm(bq_ctx* c, CommBlock* own, CommBlock* mc);
}
using namespace bq;

extern "C" {

int bq_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int bq_ctx_create(int device, bq_ctx** out) {
    if (!out) return BQ_ERR_INVALID;
    *out = nullptr;
    int ndev = bq_cuda_device_count();
    if (device < 0 || device >= ndev) {
        set_error("device %d not available (%d CUDA devices)", device, ndev);
        return BQ_ERR_INVALID;
    }
    DeviceGuard guard(device);
    bq_ctx* c = new bq_ctx();
    c->device = device;
    for (int r = 0; r < BQ_MAX_WORLD; ++r) c->peers[r] = nullptr, c->peer_ipc[r] = false;
    const int rc = ctx_init(c, device);
    if (rc != BQ_OK) {
        bq_ctx_destroy(c);      // frees whatever the failed set-up had already allocated
        return rc;
    }
    *out = c;
    return BQ_OK;
}

}  // extern "C"

namespace bq {

// Everything bq_ctx_create allocates; on any failure the caller destroys the half-built context.
static int ctx_init(bq_ctx* c, int device) {
    int sms = 0;
    BQ_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    c->sm_count = sms;
    BQ_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    BQ_CUDA(cudaEventCreate(&c->ev0));
    BQ_CUDA(cudaEventCreate(&c->ev1));
    BQ_CUDA(cudaEventCreateWithFlags(&c->ev_order, cudaEventDisableTiming));

    BQ_CUDA(cudaMalloc(&c->d_partials, sizeof(dd) * kChunks));
    BQ_CUDA(cudaMemset(c->d_partials, 0, sizeof(dd) * kChunks));
    BQ_CUDA(cudaMalloc(&c->d_ticket, 256));
    BQ_CUDA(cudaMemset(c->d_ticket, 0, 256));
    BQ_CUDA(cudaMalloc(&c->d_result, sizeof(bq_result)));
    BQ_CUDA(cudaMemset(c->d_result, 0, sizeof(bq_result)));
    BQ_CUDA(cudaMalloc(&c->d_scan_result, sizeof(bq_scan_result)));
    BQ_CUDA(cudaMemset(c->d_scan_result, 0, sizeof(bq_scan_result)));
    BQ_CUDA(cudaMalloc(&c->d_comm, sizeof(CommBlock)));
    BQ_CUDA(cudaMemset(c->d_comm, 0, sizeof(CommBlock)));
    BQ_CUDA(cudaMalloc(&c->d_table, sizeof(double) * (kMaxTable + 2)));
    BQ_CUDA(cudaMemset(c->d_table, 0, sizeof(double) * (kMaxTable + 2)));

    BQ_CUDA(cudaHostAlloc(&c->h_result, sizeof(bq_result), cudaHostAllocMapped));
    BQ_CUDA(cudaHostGetDevicePointer((void**)&c->h_result_dev, c->h_result, 0));
    BQ_CUDA(cudaHostAlloc(&c->h_scan_result, sizeof(bq_scan_result), cudaHostAllocMapped));
    BQ_CUDA(cudaHostGetDevicePointer((void**)&c->h_scan_result_dev, c->h_scan_result, 0));
    memset(c->h_result, 0, sizeof(bq_result));
    memset(c->h_scan_result, 0, sizeof(bq_scan_result));
    c->peers[0] = c->d_comm;

    int len = 0;
    const double* t = default_profile(&len);
    int rc = bq_ctx_set_profile(c, t, len);
    if (rc != BQ_OK) return rc;
    rc = scan_ctx_init(c);
    if (rc != BQ_OK) return rc;
    BQ_CUDA(cudaDeviceSynchronize());
    return BQ_OK;
}

}  // namespace bq

extern "C" {

void bq_ctx_destroy(bq_ctx* c) {
    if (!c) return;
    DeviceGuard guard(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (int r = 0; r < BQ_MAX_WORLD; ++r)
        if (c->peer_ipc[r] && c->peers[r]) cudaIpcCloseMemHandle(c->peers[r]);
    scan_ctx_free(c);
    cudaFree(c->d_partials);
    cudaFree(c->d_ticket);
    cudaFree(c->d_result);
    cudaFree(c->d_scan_result);
    if (!c->comm_external) cudaFree(c->d_comm);
    cudaFree(c->d_table);
    cudaFree(c->d_trace);
    cudaFreeHost(c->h_result);
    cudaFreeHost(c->h_scan_result);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->ev_order) cudaEventDestroy(c->ev_order);
    if (c->stream) cudaStreamDestroy(c->stream);
    cudaGetLastError();
    delete c;
}

int bq_ctx_device(const bq_ctx* c) { return c ? c->device : -1; }
void* bq_ctx_stream(const bq_ctx* c) { return c ? (void*)c->stream : nullptr; }
const bq_result* bq_ctx_device_result(const bq_ctx* c) { return c ? c->d_result : nullptr; }

int bq_ctx_set_profile(bq_ctx* c, const double* table, int len) {
    if (!c || !table || len < 2 || len > kMaxTable) {
        set_error("profile table length %d out of range [2, %d]", len, kMaxTable);
        return BQ_ERR_INVALID;
    }
    DeviceGuard guard(c->device);
    std::vector<double> padded(size_t(len) + 2, 0.0);
    memcpy(padded.data(), table, sizeof(double) * len);
    padded[len] = table[len - 1];   // pad so the bulk copy size is a multiple of 16 bytes
    BQ_CUDA(cudaStreamSynchronize(c->stream));
    BQ_CUDA(cudaMemcpy(c->d_table, padded.data(), sizeof(double) * (len + 2), cudaMemcpyHostToDevice));
    c->table_len = len;
    c->h_table.assign(table, table + len);
    return BQ_OK;
}

int bq_ctx_configure(bq_ctx* c, int ctas_per_sm, int cluster, int rows_per_tile) {
    if (!c) return BQ_ERR_INVALID;
    if (ctas_per_sm < 0 || ctas_per_sm > 8 || (cluster != 0 && cluster != 1 && cluster != 2) ||
        rows_per_tile < 0 || rows_per_tile > kMaxRows) {
        set_error("bad launch configuration");
        return BQ_ERR_INVALID;
    }
    c->ctas_per_sm = ctas_per_sm ? ctas_per_sm : 2;
    c->cluster = cluster ? cluster : 2;
    c->rows_override = uint32_t(rows_per_tile);
    return BQ_OK;
}

int bq_ctx_set_timeout_ms(bq_ctx* c, int ms) {
    if (!c || ms < 1) return BQ_ERR_INVALID;
    c->timeout_ms = uint32_t(ms);
    return BQ_OK;
}

// Per-CTA timeline of the quad kernel: (start ns, end ns, smid, tiles) per CTA.
int bq_ctx_trace(bq_ctx* c, int enable) {
    if (!c) return BQ_ERR_INVALID;
    DeviceGuard guard(c->device);
    BQ_CUDA(cudaStreamSynchronize(c->stream));
    if (enable && !c->d_trace) {
        BQ_CUDA(cudaMalloc(&c->d_trace, sizeof(uint64_t) * 4 * 4096));
        BQ_CUDA(cudaMemset(c->d_trace, 0, sizeof(uint64_t) * 4 * 4096));
    } else if (!enable && c->d_trace) {
        BQ_CUDA(cudaFree(c->d_trace));
        c->d_trace = nullptr;
    }
    return BQ_OK;
}

int bq_ctx_trace_read(bq_ctx* c, uint64_t* out, int max_ctas) {
    if (!c || !out || !c->d_trace) return 0;
    DeviceGuard guard(c->device);
    if (cudaStreamSynchronize(c->stream) != cudaSuccess) return 0;
    int n = c->trace_ctas < max_ctas ? c->trace_ctas : max_ctas;
    if (cudaMemcpy(out, c->d_trace, sizeof(uint64_t) * 4 * size_t(n), cudaMemcpyDeviceToHost) != cudaSuccess)
        return 0;
    return n;
}

int bq_ctx_inject_fault(bq_ctx* c, int kind) {
    if (!c) return BQ_ERR_INVALID;
    c->fault_flags = kind ? kFlagSkipPublish : 0u;
    return BQ_OK;
}

}  // extern "C"

namespace bq {

// The comm block of a context moves into memory someone else owns (symmetric / multicast-bound).
static int ctx_adopt_comm(bq_ctx* c, CommBlock* own, CommBlock* mc) {
    DeviceGuard guard(c->device);
    BQ_CUDA(cudaStreamSynchronize(c->stream));
    if (!c->comm_external && c->d_comm) BQ_CUDA(cudaFree(c->d_comm));
    c->d_comm = own;
    c->comm_external = true;
    c->mc_comm = mc;
    return BQ_OK;
}

int order_after_previous(bq_ctx* c, cudaStream_t stream) {
    if (!c->last_stream || c->last_stream == stream) return BQ_OK;
    cudaStreamCaptureStatus a = cudaStreamCaptureStatusNone, b = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &a) != cudaSuccess || cudaStreamIsCapturing(c->last_stream, &b) != cudaSuccess) {
        cudaGetLastError();       // e.g. the previous stream no longer exists: nothing left to order after
        return BQ_OK;
    }
    if (a != cudaStreamCaptureStatusNone || b != cudaStreamCaptureStatusNone) return BQ_OK;
    BQ_CUDA(cudaEventRecord(c->ev_order, c->last_stream));
    BQ_CUDA(cudaStreamWaitEvent(stream, c->ev_order, 0));
    return BQ_OK;
}

static LaunchConfig quad_config(const bq_ctx* c, const Plan& P) {
    LaunchConfig cfg;
    cfg.cluster = unsigned(c->cluster);
    unsigned grid = unsigned(c->sm_count * c->ctas_per_sm);
    (void)P;   // the grid is the same for every n: a cluster never owns more than a few chunks
    grid -= grid % cfg.cluster;
    if (grid < cfg.cluster) grid = cfg.cluster;
    cfg.grid = grid;
    return cfg;
}

// to_host: the kernel also writes the result block into the pinned host block (the callers that
// read it there); value_out: optional device address that receives the integral.
static int quad_enqueue(bq_ctx* c, const bq_job* job, int rank, int world, uint32_t extra_flags,
                        double* seg_out, cudaStream_t stream, bool to_host = true,
                        double* value_out = nullptr) {
    NvtxRange range("b200quad::integrate");
    // The plan travels in the kernel parameter buffer: no staging copy.
    Plan plan;
    Plan* hp = &plan;
    int rc = make_plan(*job, rank, world, c->rows_override, c->table_len, hp);
    if (rc != BQ_OK) return rc;
    hp->flags |= extra_flags | c->fault_flags;
    c->fault_flags = 0;
    hp->timeout_ms = c->timeout_ms;
    if (world > 1 && !(hp->flags & kFlagLocalOnly)) {
        for (int r = 0; r < world; ++r)
            if (!c->peers[r]) {
                set_error("rank %d: peer %d is not attached (world=%d)", rank, r, world);
                return BQ_ERR_INVALID;
            }
    }

    DeviceState st;
    memset(&st, 0, sizeof(st));
    st.partials = c->d_partials;
    st.ticket = c->d_ticket;
    st.result = c->d_result;
    st.result_host = to_host ? c->h_result_dev : nullptr;   // the kernel itself writes the block to the host
    st.value_out = value_out;
    st.seq = ++c->quad_seq;
    st.comm = c->d_comm;
    for (int r = 0; r < BQ_MAX_WORLD; ++r) st.peers[r] = c->peers[r];
    st.peers[rank] = c->d_comm;
    st.mc_comm = (c->use_nvls && world == c->world) ? c->mc_comm : nullptr;
    st.table = c->d_table;
    st.seg_out = seg_out;
    st.trace = c->d_trace;
    c->trace_ctas = c->d_trace ? int(quad_config(c, *hp).grid) : 0;
    rc = order_after_previous(c, stream);
    if (rc != BQ_OK) return rc;
    BQ_CUDA(launch_quad(*hp, st, quad_config(c, *hp), stream));
    c->launches += 1;
    c->last_stream = stream;
    return BQ_OK;
}

}  // namespace bq

extern "C" {

int bq_integrate_async(bq_ctx* c, const bq_job* job, void* stream) {
    if (!c || !job) return BQ_ERR_INVALID;
    DeviceGuard guard(c->device);
    cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
    return quad_enqueue(c, job, c->rank, c->world, 0, nullptr, s);
}

int bq_integrate_device_async(bq_ctx* c, const bq_job* job, double* d_value, void* stream) {
    if (!c || !job || !d_value) return BQ_ERR_INVALID;
    DeviceGuard guard(c->device);
    cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
    return quad_enqueue(c, job, c->rank, c->world, 0, nullptr, s, /*to_host=*/false, d_value);
}

int bq_device_result_fetch(bq_ctx* c, bq_result* out) {
    if (!c || !out) return BQ_ERR_INVALID;
    DeviceGuard guard(c->device);
    BQ_CUDA(cudaStreamSynchronize(c->last_stream ? c->last_stream : c->stream));
    BQ_CUDA(cudaMemcpy(out, c->d_result, sizeof(bq_result), cudaMemcpyDeviceToHost));
    return out->status;
}

int bq_integrate_local_async(bq_ctx* c, const bq_job* job, int rank, int world, double* dSegPairs,
                             void* stream) {
    if (!c || !job) return BQ_ERR_INVALID;
    DeviceGuard guard(c->device);
    cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
    return quad_enqueue(c, job, rank, world, kFlagLocalOnly, dSegPairs, s, /*to_host=*/false);
}

int bq_result_fetch(bq_ctx* c, bq_result* out) {
    if (!c || !out) return BQ_ERR_INVALID;
    DeviceGuard guard(c->device);
    BQ_CUDA(cudaStreamSynchronize(c->last_stream ? c->last_stream : c->stream));
    *out = *c->h_result;
    return out->status;
}

int bq_integrate(bq_ctx* c, const bq_job* job, bq_result* out) {
    if (!c || !job || !out) return BQ_ERR_INVALID;
    DeviceGuard guard(c->device);
    BQ_CUDA(cudaEventRecord(c->ev0, c->stream));
    int rc = quad_enqueue(c, job, c->rank, c->world, 0, nullptr, c->stream);
    if (rc != BQ_OK) return rc;
    BQ_CUDA(cudaEventRecord(c->ev1, c->stream));
    BQ_CUDA(cudaStreamSynchronize(c->stream));
    *out = *c->h_result;
    float ms = 0.f;
    BQ_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    out->device_ms = ms;
    if (out->status != BQ_OK) set_error("kernel reported status %d", out->status);
    return out->status;
}

int bq_integrate_value(bq_ctx* c, const bq_job* job, double* value, bq_result* out) {
    if (!c || !job) return BQ_ERR_INVALID;
    DeviceGuard guard(c->device);
    int rc = quad_enqueue(c, job, c->rank, c->world, 0, nullptr, c->stream);
    if (rc != BQ_OK) return rc;
    // The kernel writes the block's call number last (after a system fence): watch it
    // instead of paying for a stream synchronise.  Every few thousand polls ask the
    // driver, so a failed launch or a lost kernel cannot spin forever.
    const uint64_t want = c->quad_seq;
    const volatile uint64_t* flag = &c->h_result->seq;
    for (uint32_t spins = 0; *flag != want; ++spins) {
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
        if ((spins & 0xfff) == 0xfff) {
            const cudaError_t q = cudaStreamQuery(c->stream);
            if (q == cudaSuccess) break;             // finished: the block is complete
            if (q != cudaErrorNotReady) {
                set_error("integrate kernel failed: %s", cudaGetErrorString(q));
                return BQ_ERR_CUDA;
            }
        }
    }
    __atomic_thread_fence(__ATOMIC_ACQUIRE);
    if (*flag != want) BQ_CUDA(cudaStreamSynchronize(c->stream));
    const bq_result* r = c->h_result;
    if (value) *value = r->value;
    if (out) {
        *out = *r;
        out->device_ms = float(double(r->t_end_ns - r->t_start_ns) * 1e-6);   // kernel time (globaltimer)
    }
    if (r->status != BQ_OK) set_error("kernel reported status %d", r->status);
    return r->status;
}

// ---- peers ---------------------------------------------------------------

int bq_ipc_export(void* dptr, unsigned char handle[BQ_IPC_HANDLE_BYTES]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == BQ_IPC_HANDLE_BYTES, "IPC handle size");
    cudaIpcMemHandle_t h;
    BQ_CUDA(cudaIpcGetMemHandle(&h, dptr));
    memcpy(handle, &h, sizeof(h));
    return BQ_OK;
}

int bq_ipc_export_at(void* dptr, unsigned char handle[BQ_IPC_HANDLE_BYTES], uint64_t* offset) {
    if (!dptr || !offset) return BQ_ERR_INVALID;
    // A CUDA-IPC handle names the whole allocation: cudaIpcOpenMemHandle returns the allocation's
    // base on the importing side, wherever inside it dptr points (a caching allocator hands out
    // interior pointers), so the exporter ships dptr's offset from that base along with the handle.
    const drv::Api* d = drv::api();
    if (!d) {
        set_error("bq_ipc_export_at: CUDA driver entry points are not available");
        return BQ_ERR_UNSUPPORTED;
    }
    CUdeviceptr base = 0;
    size_t size = 0;
    const CUresult r = d->MemGetAddressRange(&base, &size, (CUdeviceptr)(uintptr_t)dptr);
    if (r != CUDA_SUCCESS) {
        set_error("cuMemGetAddressRange failed: %s", drv::error_string(r));
        return BQ_ERR_CUDA;
    }
    *offset = uint64_t((uintptr_t)dptr - (uintptr_t)base);
    return bq_ipc_export(dptr, handle);
}

int bq_ipc_open(const unsigned char handle[BQ_IPC_HANDLE_BYTES], void** dptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    BQ_CUDA(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
    return BQ_OK;
}

int bq_ipc_close(void* dptr) {
    BQ_CUDA(cudaIpcCloseMemHandle(dptr));
    return BQ_OK;
}

int bq_ctx_export_comm(bq_ctx* c, unsigned char handle[BQ_IPC_HANDLE_BYTES]) {
    if (!c) return BQ_ERR_INVALID;
    DeviceGuard guard(c->device);
    return bq_ipc_export(c->d_comm, handle);
}

int bq_ctx_attach_peers(bq_ctx* c, int rank, int world, const unsigned char* handles) {
    if (!c || rank < 0 || rank >= world || world > BQ_MAX_WORLD || (kSegments % world) != 0) {
        set_error("attach_peers: bad rank/world %d/%d", rank, world);
        return BQ_ERR_INVALID;
    }
    DeviceGuard guard(c->device);
    BQ_CUDA(cudaStreamSynchronize(c->last_stream ? c->last_stream : c->stream));   // no kernel may still use a mapping
    for (int r = 0; r < BQ_MAX_WORLD; ++r) {
        if (c->peer_ipc[r] && c->peers[r]) cudaIpcCloseMemHandle(c->peers[r]);
        c->peers[r] = nullptr;
        c->peer_ipc[r] = false;
    }
    for (int r = 0; r < world; ++r) {
        if (r == rank) {
            c->peers[r] = c->d_comm;
            continue;
        }
        if (!handles) return BQ_ERR_INVALID;
        void* p = nullptr;
        int rc = bq_ipc_open(handles + size_t(r) * BQ_IPC_HANDLE_BYTES, &p);
        if (rc != BQ_OK) return rc;
        c->peers[r] = (CommBlock*)p;
        c->peer_ipc[r] = true;
    }
    c->rank = rank;
    c->world = world;
    return BQ_OK;
}

int bq_ctx_adopt_comm(bq_ctx* c, int rank, int world, void* const* comm_ptrs, void* mc_ptr, size_t bytes) {
    if (!c || !comm_ptrs || rank < 0 || rank >= world || world > BQ_MAX_WORLD || (kSegments % world) != 0 ||
        bytes < sizeof(CommBlock)) {
        set_error("adopt_comm: bad rank/world %d/%d or block smaller than %zu bytes", rank, world, sizeof(CommBlock));
        return BQ_ERR_INVALID;
    }
    for (int r = 0; r < world; ++r)
        if (!comm_ptrs[r] || (uintptr_t(comm_ptrs[r]) & 255)) {
            set_error("adopt_comm: comm block of rank %d is null or not 256-byte aligned", r);
            return BQ_ERR_INVALID;
        }
    DeviceGuard guard(c->device);
    BQ_CUDA(cudaStreamSynchronize(c->last_stream ? c->last_stream : c->stream));
    for (int r = 0; r < BQ_MAX_WORLD; ++r) {
        if (c->peer_ipc[r] && c->peers[r]) cudaIpcCloseMemHandle(c->peers[r]);
        c->peers[r] = nullptr;
        c->peer_ipc[r] = false;
    }
    const int rc = ctx_adopt_comm(c, static_cast<CommBlock*>(comm_ptrs[rank]), static_cast<CommBlock*>(mc_ptr));
    if (rc != BQ_OK) return rc;
    BQ_CUDA(cudaMemset(c->d_comm, 0, sizeof(CommBlock)));
    BQ_CUDA(cudaDeviceSynchronize());
    for (int r = 0; r < world; ++r) c->peers[r] = static_cast<CommBlock*>(comm_ptrs[r]);
    c->rank = rank;
    c->world = world;
    const char* env = getenv("BQ_NVLS");
    c->use_nvls = mc_ptr != nullptr && env && env[0] == '1';
    return BQ_OK;
}

int bq_ctx_set_nvls(bq_ctx* c, int on) {
    if (!c) return BQ_ERR_INVALID;
    if (on && !c->mc_comm) {
        set_error("NVLS multicast is not available for this context");
        return BQ_ERR_UNSUPPORTED;
    }
    c->use_nvls = on != 0;
    return BQ_OK;
}

int bq_ctx_set_scan_multicast(bq_ctx* c, double* mcP, double* mcQ) {
    if (!c) return BQ_ERR_INVALID;
    c->scan_mcP = mcP;
    c->scan_mcQ = mcQ;
    return BQ_OK;
}

int bq_ctx_reset_comm(bq_ctx* c) {
    if (!c) return BQ_ERR_INVALID;
    DeviceGuard guard(c->device);
    BQ_CUDA(cudaStreamSynchronize(c->stream));
    BQ_CUDA(cudaMemset(c->d_comm, 0, sizeof(CommBlock)));
    BQ_CUDA(cudaMemset(c->d_ticket, 0, 256));
    const int rc = scan_clear_scratch(c);
    if (rc != BQ_OK) return rc;
    BQ_CUDA(cudaDeviceSynchronize());
    return BQ_OK;
}

// ---- single-process world ----------------------------------------------------

struct bq_world {
    std::vector<bq_ctx*> ctx;
    bq::SymmAlloc comm;                   // the comm blocks as one symmetric, multicast-bound allocation
    std::vector<bq::SymmAlloc> tables;    // bq_world_alloc_tables
};

int bq_world_create(int ndev, const int* devices, bq_world** out) {
    if (!out || ndev < 1 || ndev > BQ_MAX_WORLD || (kSegments % ndev) != 0) {
        set_error("world size %d must be 1, 2, 4 or 8", ndev);
        return BQ_ERR_INVALID;
    }
    bq_world* w = new bq_world();
    for (int r = 0; r < ndev; ++r) {
        bq_ctx* c = nullptr;
        int rc = bq_ctx_create(devices ? devices[r] : r, &c);
        if (rc != BQ_OK) {
            bq_world_destroy(w);
            return rc;
        }
        w->ctx.push_back(c);
    }
    // Comm blocks: one symmetric allocation (a copy per GPU, every copy mapped for every GPU) bound
    // to an NVLS multicast object when the fabric offers one; else each context keeps its own
    // cudaMalloc'd block and peers reach it through peer access.
    if (ndev > 1) {
        int devs[BQ_MAX_WORLD];
        for (int r = 0; r < ndev; ++r) devs[r] = w->ctx[r]->device;
        if (symm_create(devs, ndev, sizeof(CommBlock), /*want_mc=*/true, &w->comm) == BQ_OK) {
            const char* env = getenv("BQ_NVLS");
            const bool nvls = w->comm.mc != nullptr && env && env[0] == '1';
            for (int r = 0; r < ndev; ++r) {
                if (ctx_adopt_comm(w->ctx[r], (CommBlock*)w->comm.uc[r], (CommBlock*)w->comm.mc) != BQ_OK) {
                    bq_world_destroy(w);
                    return BQ_ERR_CUDA;
                }
                w->ctx[r]->use_nvls = nvls;
            }
        }
    }
    for (int r = 0; r < ndev; ++r) {
        bq_ctx* c = w->ctx[r];
        DeviceGuard guard(c->device);
        for (int q = 0; q < ndev; ++q) {
            if (q != r && w->ctx[q]->device != c->device) {
                int can = 0;
                cudaDeviceCanAccessPeer(&can, c->device, w->ctx[q]->device);
                if (!can) {
                    set_error("device %d cannot access device %d", c->device, w->ctx[q]->device);
                    bq_world_destroy(w);
                    return BQ_ERR_UNSUPPORTED;
                }
                cudaError_t e = cudaDeviceEnablePeerAccess(w->ctx[q]->device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                    set_error("cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
                    bq_world_destroy(w);
                    return BQ_ERR_CUDA;
                }
                cudaGetLastError();
            }
            c->peers[q] = w->ctx[q]->d_comm;
        }
        c->rank = r;
        c->world = ndev;
    }
    *out = w;
    return BQ_OK;
}

void bq_world_destroy(bq_world* w) {
    if (!w) return;
    for (bq_ctx* c : w->ctx) bq_ctx_destroy(c);     // synchronises every stream first
    for (SymmAlloc& t : w->tables) symm_destroy(&t);
    symm_destroy(&w->comm);
    delete w;
}

int bq_world_has_multicast(const bq_world* w) { return (w && w->comm.mc) ? 1 : 0; }

int bq_world_set_nvls(bq_world* w, int on) {
    if (!w) return BQ_ERR_INVALID;
    if (on && !w->comm.mc) {
        set_error("NVLS multicast is not available for this world (driver / fabric)");
        return BQ_ERR_UNSUPPORTED;
    }
    for (bq_ctx* c : w->ctx) c->use_nvls = on != 0;
    return BQ_OK;
}

int bq_world_alloc_tables(bq_world* w, uint64_t n, double** P, double** Q) {
    if (!w || !P || !Q || n < 1) return BQ_ERR_INVALID;
    const int nd = int(w->ctx.size());
    int devs[BQ_MAX_WORLD];
    for (int r = 0; r < nd; ++r) devs[r] = w->ctx[r]->device;
    for (int k = 0; k < 2; ++k) {
        SymmAlloc a;
        const int rc = symm_create(devs, nd, size_t(n) * sizeof(double), /*want_mc=*/true, &a);
        if (rc != BQ_OK) return rc;
        w->tables.push_back(a);
        for (int r = 0; r < nd; ++r) (k == 0 ? P : Q)[r] = static_cast<double*>(a.uc[r]);
    }
    return BQ_OK;
}

void bq_world_free_tables(bq_world* w) {
    if (!w) return;
    for (bq_ctx* c : w->ctx) {
        DeviceGuard guard(c->device);
        cudaStreamSynchronize(c->last_stream ? c->last_stream : c->stream);
        c->scan_mcP = c->scan_mcQ = nullptr;
    }
    for (SymmAlloc& t : w->tables) symm_destroy(&t);
    w->tables.clear();
}

// Multicast address of a table handed out by bq_world_alloc_tables (null for any other pointer).
double* bq_world_table_multicast(bq_world* w, const double* rank0_ptr) {
    if (!w) return nullptr;
    for (SymmAlloc& t : w->tables)
        if (t.uc[0] == rank0_ptr) return static_cast<double*>(t.mc);
    return nullptr;
}

int bq_world_size(const bq_world* w) { return w ? int(w->ctx.size()) : 0; }
bq_ctx* bq_world_ctx(bq_world* w, int rank) {
    return (w && rank >= 0 && rank < int(w->ctx.size())) ? w->ctx[rank] : nullptr;
}

int bq_world_integrate(bq_world* w, const bq_job* job, bq_result* out) {
    if (!w || !job) return BQ_ERR_INVALID;
    const int n = int(w->ctx.size());
    // Everything that can be rejected is checked for EVERY rank before the first launch: a rank
    // launched alone would sit at the in-kernel barrier until its time-out.
    for (int r = 0; r < n; ++r) {
        Plan probe;
        const int rc = make_plan(*job, r, n, w->ctx[r]->rows_override, w->ctx[r]->table_len, &probe);
        if (rc != BQ_OK) return rc;
        for (int q = 0; q < n; ++q)
            if (!w->ctx[r]->peers[q]) {
                set_error("rank %d: peer %d is not attached (world=%d)", r, q, n);
                return BQ_ERR_INVALID;
            }
    }
    // Enqueue everywhere first: the kernels meet at the in-kernel barrier, so
    // no rank may be waited for before all have been launched.
    for (int r = 0; r < n; ++r) {
        bq_ctx* c = w->ctx[r];
        DeviceGuard guard(c->device);
        int rc = cudaEventRecord(c->ev0, c->stream) == cudaSuccess ? BQ_OK : BQ_ERR_CUDA;
        if (rc == BQ_OK) rc = quad_enqueue(c, job, r, n, 0, nullptr, c->stream);
        if (rc == BQ_OK && cudaEventRecord(c->ev1, c->stream) != cudaSuccess) rc = BQ_ERR_CUDA;
        if (rc != BQ_OK) {
            // A launch failed after others went out: let those run into their time-out, then put
            // every barrier word of the world back to zero so the next call starts clean.
            for (int q = 0; q < r; ++q) {
                DeviceGuard g2(w->ctx[q]->device);
                cudaStreamSynchronize(w->ctx[q]->stream);
            }
            for (int q = 0; q < n; ++q) bq_ctx_reset_comm(w->ctx[q]);
            return rc;
        }
    }
    int worst = BQ_OK;
    for (int r = 0; r < n; ++r) {
        bq_ctx* c = w->ctx[r];
        DeviceGuard guard(c->device);
        BQ_CUDA(cudaStreamSynchronize(c->stream));
        float ms = 0.f;
        BQ_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
        c->h_result->device_ms = ms;
        if (out) out[r] = *c->h_result;
        if (c->h_result->status != BQ_OK) worst = c->h_result->status;
    }
    return worst;
}

}  // extern "C"

// ---- process-wide defaults ------------------------------------------------------
// One context per device and one world per size, created on first use and shared by everything
// in the process that does not bring its own: the Python layer (parallel.context(),
// parallel.local_world()) and the torch ops resolve to the SAME objects, so a table set through one
// is the table the other integrates, and their calls are ordered against each other.
namespace {
std::mutex g_default_mu;
bq_ctx* g_default_ctx[64] = {};
bq_world* g_default_world[BQ_MAX_WORLD + 1] = {};
bq_ctx* g_dist_ctx = nullptr;
}  // namespace

extern "C" {

bq_ctx* bq_default_ctx(int device) {
    if (device < 0 || device >= 64) {
        set_error("bq_default_ctx: device %d out of range", device);
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_default_mu);
    if (!g_default_ctx[device] && bq_ctx_create(device, &g_default_ctx[device]) != BQ_OK) g_default_ctx[device] = nullptr;
    return g_default_ctx[device];
}

bq_world* bq_default_world(int size) {
    if (size < 1 || size > BQ_MAX_WORLD) {
        set_error("bq_default_world: size %d out of range", size);
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_default_mu);
    if (!g_default_world[size] && bq_world_create(size, nullptr, &g_default_world[size]) != BQ_OK) g_default_world[size] = nullptr;
    return g_default_world[size];
}

// The context of this process's rank in the torch.distributed world (registered by DistWorld).
int bq_set_dist_ctx(bq_ctx* c) {
    std::lock_guard<std::mutex> lock(g_default_mu);
    g_dist_ctx = c;
    return BQ_OK;
}
bq_ctx* bq_dist_ctx(void) {
    std::lock_guard<std::mutex> lock(g_default_mu);
    return g_dist_ctx;
}
int bq_ctx_world(const bq_ctx* c) { return c ? c->world : 0; }
int bq_ctx_rank(const bq_ctx* c) { return c ? c->rank : -1; }

}  // extern "C"

// ---- small device-memory helpers for C callers (the CLIs) -------------------
extern "C" {

int bq_device_alloc(int device, size_t bytes, void** out) {
    if (!out) return BQ_ERR_INVALID;
    DeviceGuard guard(device);
    BQ_CUDA(cudaMalloc(out, bytes));
    return BQ_OK;
}

int bq_device_free(int device, void* p) {
    DeviceGuard guard(device);
    BQ_CUDA(cudaFree(p));
    return BQ_OK;
}

int bq_device_read(int device, void* host_dst, const void* dev_src, size_t bytes) {
    DeviceGuard guard(device);
    BQ_CUDA(cudaMemcpy(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost));
    return BQ_OK;
}

}  // extern "C"
