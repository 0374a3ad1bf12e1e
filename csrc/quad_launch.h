// Host-side launch interface of the fused kernels.
#pragma once
#include <cuda_runtime.h>

#include "device_state.h"
#include "plan.h"

namespace bq {

struct LaunchConfig {
    unsigned grid;      // CTAs, a multiple of `cluster`
    unsigned cluster;   // 1 or 2
};

cudaError_t launch_quad(const Plan& P, const DeviceState& st, const LaunchConfig& cfg,
                        cudaStream_t stream);

}  // namespace bq
