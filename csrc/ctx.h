// Definition of the opaque bq_ctx (one per process x device).
#pragma once
#include <cuda_runtime.h>

#include <vector>

#include "device_state.h"
#include "plan.h"
#include "scan_plan.h"

namespace bq {

}  // namespace bq

struct bq_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t last_stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t ev_order = nullptr;       // orders a launch after the previous one when the stream changes

    // launch knobs
    int ctas_per_sm = 2;
    int cluster = 2;
    uint32_t rows_override = 0;
    uint32_t timeout_ms = 2000;
    uint32_t fault_flags = 0;

    // device buffers (allocated once)
    bq::dd* d_partials = nullptr;
    unsigned int* d_ticket = nullptr;     // [0] quad, [8] scan phase A, [16] scan phase B
    bq_result* d_result = nullptr;
    bq_scan_result* d_scan_result = nullptr;
    bq::CommBlock* d_comm = nullptr;
    bool comm_external = false;           // d_comm lives in symmetric memory owned by the world / the caller
    bq::CommBlock* mc_comm = nullptr;     // NVLS multicast address of every rank's comm block (or null)
    bool use_nvls = false;                // publish through mc_comm (multimem.st) instead of per-peer stores
    double* d_table = nullptr;
    uint64_t* d_trace = nullptr;          // per-CTA timeline, allocated by bq_ctx_trace(on)
    int trace_ctas = 0;
    int table_len = 0;
    std::vector<double> h_table;

    // scan scratch: LL descriptors (grown on demand, outside the timed path; see scan_launch.h)
    unsigned char* d_scan_scratch = nullptr;   // tile aggregates | group totals | supergroup prefixes
    uint64_t* d_scan_misc = nullptr;      // [0..63] segment-base descriptors, [64] abort word
    uint64_t scan_tiles_cap = 0;
    double* scan_peerP[BQ_MAX_WORLD] = {};
    double* scan_peerQ[BQ_MAX_WORLD] = {};
    double* scan_mcP = nullptr;           // multicast addresses of the replicated tables (or null)
    double* scan_mcQ = nullptr;
    int scan_grid = 0;

    // pinned staging
    bq_result* h_result = nullptr;
    bq_result* h_result_dev = nullptr;    // h_result as the device addresses it (mapped)
    bq_scan_result* h_scan_result = nullptr;
    bq_scan_result* h_scan_result_dev = nullptr;

    // world
    int rank = 0, world = 1;
    bq::CommBlock* peers[BQ_MAX_WORLD];
    bool peer_ipc[BQ_MAX_WORLD];

    uint64_t launches = 0;
    uint64_t quad_seq = 0;                // call number of the latest integrate launch
};

namespace bq {
int scan_ctx_init(bq_ctx* c);
void scan_ctx_free(bq_ctx* c);
int scan_clear_scratch(bq_ctx* c);
// One context = one set of scratch (chunk sums, tickets, result blocks, barrier epochs): two of its
// kernels must never overlap on the GPU.  Calls on one stream are ordered by the stream; when a call
// arrives on a DIFFERENT stream than the previous one, this makes the new stream wait for everything
// enqueued so far on the old one (no-op while either stream is being captured: a captured launch is
// ordered by the graph it ends up in).
int order_after_previous(bq_ctx* c, cudaStream_t stream);
}  // namespace bq
