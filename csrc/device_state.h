// Device-side state shared between the host runtime (ctx.cu) and the kernels.
#pragma once
#include <stdint.h>

#include "bq_dd.cuh"
#include "plan.h"

namespace bq {

// One per GPU, allocated with cudaMalloc so it can be exported through CUDA IPC
// and mapped by every peer.  Peers only ever write `slots`, `scan_slots` and
// add to the two `arrived` counters; `epoch` words are private to the owner.
struct alignas(256) CommBlock {
    uint64_t arrived;         // quad barrier: += 1 per rank per call
    uint64_t scan_arrived;    // scan barrier 1 (segment totals published)
    uint64_t scan_done;       // scan barrier 2 (tables + final scalars stored)
    uint64_t epoch;           // calls completed by the owner (quad)
    uint64_t scan_epoch;      // calls completed by the owner (scan)
    uint64_t pad[3];
    double slots[2][kSegments][2];            // [parity][segment](hi, lo)
    double scan_slots[2][kSegments][4];       // [parity][segment](P_hi, P_lo, Q_hi, Q_lo)
    double scan_final[4];                     // P[n-1], Q[n-1], P[n-2], spare
};

struct DeviceState {
    dd* partials;             // [kChunks] chunk sums
    unsigned int* ticket;     // clusters finished (reset by the last one)
    bq_result* result;        // device result block
    bq_result* result_host;   // pinned host result block as mapped on this device (or null)
    double* value_out;        // optional: the integral also goes here (device memory, e.g. a tensor)
    uint64_t seq;             // call number: written last into the host block (completion word)
    CommBlock* comm;          // own comm block
    CommBlock* peers[BQ_MAX_WORLD];   // comm block of rank r as mapped here
    const double* table;      // PROFILE table (padded to an even count)
    double* seg_out;          // kFlagLocalOnly: 2*kSegments doubles, else null
    uint64_t* trace;          // optional per-CTA timeline: [cta](start ns, end ns, smid, tiles)
};

}  // namespace bq
