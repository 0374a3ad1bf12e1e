#include "drv.h"

#include <mutex>

namespace bq {
namespace drv {

namespace {
Api g_api;
std::once_flag g_once;

template <class Fn>
bool resolve(const char* name, Fn*& out) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
        cudaGetLastError();
        return false;
    }
    out = reinterpret_cast<Fn*>(p);
    return true;
}
}  // namespace

const Api* api() {
    std::call_once(g_once, [] {
        Api a;
        bool ok = true;
        ok &= resolve("cuGetErrorString", a.GetErrorString);
        ok &= resolve("cuDeviceGet", a.DeviceGet);
        ok &= resolve("cuDeviceGetAttribute", a.DeviceGetAttribute);
        ok &= resolve("cuMemGetAddressRange", a.MemGetAddressRange);
        ok &= resolve("cuMemGetAllocationGranularity", a.MemGetAllocationGranularity);
        ok &= resolve("cuMemCreate", a.MemCreate);
        ok &= resolve("cuMemRelease", a.MemRelease);
        ok &= resolve("cuMemAddressReserve", a.MemAddressReserve);
        ok &= resolve("cuMemAddressFree", a.MemAddressFree);
        ok &= resolve("cuMemMap", a.MemMap);
        ok &= resolve("cuMemUnmap", a.MemUnmap);
        ok &= resolve("cuMemSetAccess", a.MemSetAccess);
        ok &= resolve("cuMulticastCreate", a.MulticastCreate);
        ok &= resolve("cuMulticastAddDevice", a.MulticastAddDevice);
        ok &= resolve("cuMulticastBindMem", a.MulticastBindMem);
        ok &= resolve("cuMulticastGetGranularity", a.MulticastGetGranularity);
        a.ok = ok;
        g_api = a;
    });
    return g_api.ok ? &g_api : nullptr;
}

const char* error_string(CUresult r) {
    const Api* a = api();
    const char* s = nullptr;
    if (a && a->GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
    return "unknown driver error";
}

}  // namespace drv
}  // namespace bq
