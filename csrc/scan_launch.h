// Host-side launch interface of the fused scan kernel.
#pragma once
#include <cuda_runtime.h>

#include "device_state.h"
#include "scan_plan.h"

namespace bq {

// Scratch of one scan call.  Everything in it is an "LL" descriptor: 8-byte words that carry 32
// bits of payload and the 32-bit generation of the launch that wrote them, so a word is valid
// exactly when its upper half equals the running launch's generation.  Nothing is ever cleared
// between calls, no fence or flag accompanies a descriptor, and readers poll the words themselves.
struct ScanState {
    uint64_t* a_ll;           // [rank tiles][4]        tile aggregate (sum v, sum of local running sums)
    uint64_t* aex_ll;         // [rank tiles][8]        Ascan(g)[k-1]: prefix of the tile inside its group (dd P, dd Q)
    uint64_t* gt_ll;          // [rank groups][8]       group total
    uint64_t* head_ll;        // [rank groups][8]       SGP(sg-1) (+) GTscan(sg)[gi-1]: prefix of the group inside its segment
    uint64_t* sgp_ll;         // [rank supergroups][8]  in-segment prefix at the end of a supergroup
    uint64_t* segbase_ll;     // [kSegments][8]         exclusive prefix of each segment
    uint64_t* abort_word;     // a warp that timed out stores the generation here; everyone else stops waiting
    unsigned int* ticket;     // CTAs finished (reset by the last one)
    bq_scan_result* result;
    bq_scan_result* result_host;   // pinned host block as mapped on this device (or null)
    CommBlock* comm;          // own comm block: generation word, segment totals, final scalars
    CommBlock* peers[BQ_MAX_WORLD];
    CommBlock* mc_comm;       // multicast address of all ranks' comm blocks (NVLS) or null
    const double* table;
    double* P;                // this rank's view of the output tables
    double* Q;
    double* samples;
    double* peerP[BQ_MAX_WORLD];   // replicate: every rank's tables as mapped here
    double* peerQ[BQ_MAX_WORLD];
    double* mcP;              // replicate over NVLS: multicast addresses of the tables (or null)
    double* mcQ;
};

// Bytes of scratch a rank needs for `tiles` tiles (the five descriptor arrays, 256-byte aligned parts).
struct ScanScratchLayout {
    size_t a_off, aex_off, gt_off, head_off, sgp_off, total;
};
constexpr size_t scan_up256(size_t v) { return (v + 255) & ~size_t(255); }
constexpr ScanScratchLayout scan_scratch_layout(uint64_t tiles) {
    const uint64_t groups = tiles / kScanGroup + kSegments + 1;          // every segment may end in a partial group
    const uint64_t supers = groups / kScanSuper + kSegments + 1;
    ScanScratchLayout L{};
    L.a_off = 0;
    L.aex_off = scan_up256(size_t(tiles) * 4 * sizeof(uint64_t));
    L.gt_off = L.aex_off + scan_up256(size_t(tiles) * 8 * sizeof(uint64_t));
    L.head_off = L.gt_off + scan_up256(size_t(groups) * 8 * sizeof(uint64_t));
    L.sgp_off = L.head_off + scan_up256(size_t(groups) * 8 * sizeof(uint64_t));
    L.total = L.sgp_off + scan_up256(size_t(supers) * 8 * sizeof(uint64_t));
    return L;
}

int scan_max_grid(int device, int sm_count);
cudaError_t launch_scan(const ScanPlan& P, const ScanState& st, unsigned grid, cudaStream_t stream);

}  // namespace bq
