// Device-side state shared between the host runtime (ctx.cu) and the kernels.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include "bq_dd.cuh"
#include "plan.h"

namespace bq {

// One per GPU, in peer-visible memory (cudaMalloc + CUDA IPC, or a cuMem allocation bound to a
// multicast object): every peer maps it.  Peers only ever write the `*_ll` words; the `epoch` words
// are private to the owner.
//
// The partial results cross NVLink as self-validating words ("LL" layout): every 8-byte word
// carries 32 bits of payload and the 32-bit call number, so one plain store per word is the
// whole message — no fence, no separate flag, no counter: the reader polls the words themselves
// and a word whose upper half equals the call number is complete (8-byte stores are
// single-copy atomic).  A dd pair is 4 such words = one 32-byte store.
struct alignas(256) CommBlock {
    uint64_t epoch;           // calls completed by the owner (quad)
    uint64_t scan_epoch;      // scan launches completed by the owner = generation of its descriptors
    uint64_t pad[6];
    uint64_t ll[2][kSegments][4];             // quad  [parity][segment]: hi.lo32, hi.hi32, lo.lo32, lo.hi32 | call << 32
    uint64_t scan_ll[2][kSegments][8];        // scan  [parity][segment]: segment total (P_hi, P_lo, Q_hi, Q_lo), 2 words each
    uint64_t scan_final_ll[2][8];             // scan  [parity]: P[n-1], Q[n-1], P[n-2] (2 words each), 2 spare
    uint64_t scan_done_ll[2][BQ_MAX_WORLD];   // scan  [parity][rank]: "my stores into your tables are complete" (replicate)
};
static_assert(offsetof(CommBlock, ll) % 32 == 0 && offsetof(CommBlock, scan_ll) % 32 == 0, "LL rows are 16-byte aligned runs");

struct DeviceState {
    dd* partials;             // [kChunks] chunk sums
    unsigned int* ticket;     // clusters finished (reset by the last one)
    bq_result* result;        // device result block
    bq_result* result_host;   // pinned host result block as mapped on this device (or null)
    double* value_out;        // optional: the integral also goes here (device memory, e.g. a tensor)
    uint64_t seq;             // call number: written last into the host block (completion word)
    CommBlock* comm;          // own comm block
    CommBlock* peers[BQ_MAX_WORLD];   // comm block of rank r as mapped here
    CommBlock* mc_comm;       // multicast (NVLS) address of all ranks' comm blocks, or null: one
                              // multimem.st then reaches every rank's copy through the switch
    const double* table;      // PROFILE table (padded to an even count)
    double* seg_out;          // kFlagLocalOnly: 2*kSegments doubles, else null
    uint64_t* trace;          // optional per-CTA timeline: [cta](start ns, end ns, smid, tiles)
};

}  // namespace bq
