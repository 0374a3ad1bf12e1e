// Host-side launch interface of the fused scan kernel.
#pragma once
#include <cuda_runtime.h>

#include "device_state.h"
#include "scan_plan.h"

namespace bq {

struct ScanState {
    double* agg;              // [rank tiles][2]  tile aggregates (sum v, sum of local P)
    double* prefix;           // [rank tiles][4]  exclusive prefix inside the segment (dd P, dd Q)
    double* segbase;          // [kSegments][4]   exclusive prefix of each segment (dd P, dd Q)
    uint64_t* ready;          // phase A -> B release word (epoch valued)
    unsigned int* ticket_a;
    unsigned int* ticket_b;
    bq_scan_result* result;
    bq_scan_result* result_host;   // pinned host block as mapped on this device (or null)
    CommBlock* comm;
    CommBlock* peers[BQ_MAX_WORLD];
    const double* table;
    double* P;                // this rank's view of the output tables
    double* Q;
    double* samples;
    double* peerP[BQ_MAX_WORLD];   // replicate: every rank's tables as mapped here
    double* peerQ[BQ_MAX_WORLD];
};

int scan_max_grid(int device, int sm_count);
cudaError_t launch_scan(const ScanPlan& P, const ScanState& st, unsigned grid, cudaStream_t stream);

}  // namespace bq
