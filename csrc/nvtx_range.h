// NVTX ranges around the host entry points (header-only NVTX 3: no link-time
// dependency, a no-op costing a few nanoseconds unless a profiler is attached).
// The reference has no tracing beyond clock_gettime around main()
// (riemann.cpp:51,91; 4main.c:67,238; cintegrate.cu:104,139).
#pragma once
#include <nvtx3/nvToolsExt.h>

namespace bq {

struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

}  // namespace bq
