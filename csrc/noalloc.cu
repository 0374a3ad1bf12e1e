// bq_cumulative_noalloc: cumulative integration for callers that must not allocate.
//
// Who needs this: code that runs where device allocation is off the table — a
// real-time or latency-critical loop that forbids cudaMalloc/cudaFree because
// they synchronise the device and may take milliseconds; a plug-in living
// inside a host application that owns all device memory and hands out only
// fixed input/output buffers; a process running close to the memory limit
// where an allocation can simply fail.  The regular entry points
// (bq_cumulative, bq_ctx_create) allocate a context, pinned result blocks and
// scratch that grows with n; this one allocates nothing, ever.
//
// How: the scratch of the fused scan kernel (the descriptors of its tile
// aggregates, group totals and supergroup prefixes, segment bases, the ticket,
// the generation word, the result block and a copy of the velocity table) lives
// in ONE file-scope `static __device__ __align__(256)` array below.  Its size is
// fixed at build time, which is what bounds n (BQ_NOALLOC_MAX_SAMPLES).  Every
// GPU that loads this library carries the array (about 3.3 MB) until the process
// exits.  It always integrates the library's default velocity table: a table set
// with bq_ctx_set_profile belongs to a context, and this entry point has none.
//
// Rules a caller has to know (also in include/b200quad.h):
//  * The first call on a device performs the one-time set-up for that device:
//    it resolves the array's device address with cudaGetSymbolAddress (in host
//    code the array's name is only a host-side handle, never a device
//    pointer), zeroes it and uploads the table.  That set-up uses synchronous
//    runtime calls, so make the first call on each device before starting any
//    stream capture.
//  * All calls on a device share the one array.  Two calls that overlap on the
//    GPU (different streams, or the same call issued again before the first
//    finished on another stream) race on it; serialise them on one stream.
//  * The array keeps whatever the previous call left.  What must start at zero
//    (the ticket, the abort word, the result block) is cleared by a
//    stream-ordered memset at the start of every call.  The descriptor areas
//    need no clearing: a descriptor word is valid only when it carries the
//    generation of the running launch, and the generation word — which the
//    memset deliberately leaves alone — moves on with every call.
//  * Single GPU only (a world of 1): peers would need mapped memory.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include <mutex>

#include "host_common.h"
#include "scan_launch.h"

namespace bq {
namespace {

constexpr uint64_t kNoAllocMaxTiles = BQ_NOALLOC_MAX_SAMPLES / kScanTile;
static_assert(kNoAllocMaxTiles * kScanTile == BQ_NOALLOC_MAX_SAMPLES, "limit must be whole tiles");
constexpr int kMaxDevices = 64;

// Layout of the static scratch.  `control` is what a call clears up front.
struct NoAllocControl {
    bq_scan_result result;
    uint64_t abort_word;
    unsigned int ticket[8];
};

constexpr size_t kNoAllocDescBytes = scan_scratch_layout(kNoAllocMaxTiles).total;

struct NoAllocScratch {
    NoAllocControl control;
    CommBlock comm;                               // generation word, segment totals, final scalars (never cleared)
    alignas(32) uint64_t segbase_ll[8 * kSegments];
    alignas(32) double table[kMaxTable + 2];      // PROFILE table (default ex4vel.h values)
    alignas(256) unsigned char desc[kNoAllocDescBytes];   // tile / group / supergroup descriptors
};

// The scratch itself.  Host code must never pass this name to a kernel: see set-up below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

struct DeviceSetup {
    bool ready = false;
    NoAllocScratch* scratch = nullptr;   // device address from cudaGetSymbolAddress
    int grid = 0;
    int table_len = 0;
};

std::mutex g_mu;
DeviceSetup g_setup[kMaxDevices];

#define BQ_NA_CUDA(expr)                                                                     \
    do {                                                                                     \
        cudaError_t e_ = (expr);                                                             \
        if (e_ != cudaSuccess) {                                                             \
            set_error("bq_cumulative_noalloc: %s failed: %s", #expr, cudaGetErrorString(e_)); \
            return BQ_ERR_CUDA;                                                              \
        }                                                                                    \
    } while (0)

// One-time, per device, with that device current.
int setup_device(int device, DeviceSetup** out) {
    std::lock_guard<std::mutex> lock(g_mu);
    DeviceSetup& s = g_setup[device];
    if (!s.ready) {
        void* ptr = nullptr;
        BQ_NA_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        NoAllocScratch* scratch = static_cast<NoAllocScratch*>(ptr);
        BQ_NA_CUDA(cudaMemset(scratch, 0, sizeof(NoAllocScratch)));
        int len = 0;
        const double* table = default_profile(&len);
        BQ_NA_CUDA(cudaMemcpy(scratch->table, table, sizeof(double) * size_t(len), cudaMemcpyHostToDevice));
        // pad entry so the kernel's bulk copy (a multiple of 16 bytes) reads defined data
        BQ_NA_CUDA(cudaMemcpy(scratch->table + len, table + len - 1, sizeof(double), cudaMemcpyHostToDevice));
        int sms = 0;
        BQ_NA_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        s.grid = scan_max_grid(device, sms);
        s.table_len = len;
        s.scratch = scratch;
        BQ_NA_CUDA(cudaDeviceSynchronize());
        s.ready = true;
    }
    *out = &s;
    return BQ_OK;
}

}  // namespace
}  // namespace bq

using namespace bq;

extern "C" {

uint64_t bq_noalloc_max_samples(void) { return BQ_NOALLOC_MAX_SAMPLES; }

int bq_cumulative_noalloc(int device, const bq_scan_job* job, double* dP, double* dQ,
                          bq_scan_result* out, void* stream) {
    if (!job || !dP || !dQ) {
        set_error("bq_cumulative_noalloc: job, P and Q are required");
        return BQ_ERR_INVALID;
    }
    if ((uintptr_t(dP) | uintptr_t(dQ)) & 31) {
        set_error("bq_cumulative_noalloc: P and Q must be 32-byte aligned");
        return BQ_ERR_INVALID;
    }
    if (job->n > BQ_NOALLOC_MAX_SAMPLES) {
        set_error("bq_cumulative_noalloc: n=%llu exceeds the no-allocation limit of %llu samples "
                  "(fixed static scratch); use bq_cumulative for larger grids",
                  (unsigned long long)job->n, (unsigned long long)BQ_NOALLOC_MAX_SAMPLES);
        return BQ_ERR_INVALID;
    }
    if (device < 0 || device >= kMaxDevices) {
        set_error("bq_cumulative_noalloc: device %d not available", device);
        return BQ_ERR_INVALID;
    }
    // Hot path of a latency-critical caller: no device-count query per call; a device that does
    // not exist is reported by cudaSetDevice / the set-up below.
    int prev = -1;
    BQ_NA_CUDA(cudaGetDevice(&prev));
    if (prev != device && cudaSetDevice(device) != cudaSuccess) {
        cudaGetLastError();
        set_error("bq_cumulative_noalloc: device %d not available", device);
        return BQ_ERR_INVALID;
    }
    struct Restore {
        int prev, device;
        ~Restore() {
            if (prev >= 0 && prev != device) cudaSetDevice(prev);
        }
    } restore{prev, device};

    DeviceSetup* s = nullptr;
    int rc = setup_device(device, &s);
    if (rc != BQ_OK) return rc;

    ScanPlan plan;
    rc = make_scan_plan(*job, /*rank=*/0, /*world=*/1, s->table_len, &plan);
    if (rc != BQ_OK) return rc;
    // world of 1: nothing to replicate; samples are not stored by this entry point.  The evaluation
    // flag stays: it selects the same sin/cos as bq_cumulative, so the tables are equal bit for bit.
    plan.flags &= kScanFastTrig;
    if (plan.n_tiles > kNoAllocMaxTiles) {
        set_error("bq_cumulative_noalloc: n=%llu needs %llu tiles, the static scratch holds %llu",
                  (unsigned long long)plan.n, (unsigned long long)plan.n_tiles,
                  (unsigned long long)kNoAllocMaxTiles);
        return BQ_ERR_INVALID;
    }

    cudaStream_t st = static_cast<cudaStream_t>(stream);
    NoAllocScratch* sc = s->scratch;
    // The shared array still holds the previous call's state: zero what must start at zero.
    BQ_NA_CUDA(cudaMemsetAsync(&sc->control, 0, sizeof(NoAllocControl), st));
    // (The parameter block itself travels in the kernel's launch parameters.)

    ScanState state;
    memset(&state, 0, sizeof(state));
    constexpr ScanScratchLayout L = scan_scratch_layout(kNoAllocMaxTiles);
    state.a_ll = reinterpret_cast<uint64_t*>(sc->desc + L.a_off);
    state.aex_ll = reinterpret_cast<uint64_t*>(sc->desc + L.aex_off);
    state.gt_ll = reinterpret_cast<uint64_t*>(sc->desc + L.gt_off);
    state.head_ll = reinterpret_cast<uint64_t*>(sc->desc + L.head_off);
    state.sgp_ll = reinterpret_cast<uint64_t*>(sc->desc + L.sgp_off);
    state.segbase_ll = sc->segbase_ll;
    state.abort_word = &sc->control.abort_word;
    state.ticket = sc->control.ticket;
    state.result = &sc->control.result;
    state.comm = &sc->comm;
    state.peers[0] = &sc->comm;
    state.table = sc->table;
    state.P = dP;
    state.Q = dQ;
    state.samples = nullptr;
    unsigned grid = unsigned(s->grid);
    if (plan.n_tiles < grid) grid = unsigned(plan.n_tiles ? plan.n_tiles : 1);
    BQ_NA_CUDA(launch_scan(plan, state, grid, st));
    if (out)
        BQ_NA_CUDA(cudaMemcpyAsync(out, &sc->control.result, sizeof(bq_scan_result), cudaMemcpyDefault, st));
    return BQ_OK;
}

}  // extern "C"
