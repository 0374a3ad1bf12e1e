#include "symm.h"

#include <cuda_runtime.h>
#include <string.h>

#include "drv.h"
#include "host_common.h"

namespace bq {

namespace {
#define BQ_DRV(expr)                                                                    \
    do {                                                                                \
        CUresult r_ = (expr);                                                           \
        if (r_ != CUDA_SUCCESS) {                                                       \
            set_error("%s failed: %s (%s:%d)", #expr, drv::error_string(r_), __FILE__, __LINE__); \
            return BQ_ERR_CUDA;                                                         \
        }                                                                               \
    } while (0)

size_t round_up(size_t v, size_t g) { return (v + g - 1) / g * g; }

int map_for_all(const drv::Api* d, CUmemGenericAllocationHandle h, size_t bytes, size_t gran, const int* devices,
                int n, void** out) {
    CUdeviceptr va = 0;
    BQ_DRV(d->MemAddressReserve(&va, bytes, gran, 0, 0));
    CUresult r = d->MemMap(va, bytes, 0, h, 0);
    if (r != CUDA_SUCCESS) {
        d->MemAddressFree(va, bytes);
        set_error("cuMemMap failed: %s", drv::error_string(r));
        return BQ_ERR_CUDA;
    }
    CUmemAccessDesc acc[BQ_MAX_WORLD];
    for (int i = 0; i < n; ++i) {
        acc[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc[i].location.id = devices[i];
        acc[i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    }
    r = d->MemSetAccess(va, bytes, acc, size_t(n));
    if (r != CUDA_SUCCESS) {
        d->MemUnmap(va, bytes);
        d->MemAddressFree(va, bytes);
        set_error("cuMemSetAccess failed: %s", drv::error_string(r));
        return BQ_ERR_CUDA;
    }
    *out = reinterpret_cast<void*>(va);
    return BQ_OK;
}
}  // namespace

bool symm_multicast_supported(const int* devices, int n) {
    const drv::Api* d = drv::api();
    if (!d) return false;
    for (int i = 0; i < n; ++i) {
        CUdevice dev;
        int ok = 0;
        if (d->DeviceGet(&dev, devices[i]) != CUDA_SUCCESS) return false;
        if (d->DeviceGetAttribute(&ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS || !ok) return false;
    }
    return true;
}

void symm_destroy(SymmAlloc* a) {
    const drv::Api* d = drv::api();
    if (!a || !d) return;
    if (a->mc) {
        d->MemUnmap((CUdeviceptr)(uintptr_t)a->mc, a->bytes);
        d->MemAddressFree((CUdeviceptr)(uintptr_t)a->mc, a->bytes);
    }
    for (int i = 0; i < a->n; ++i) {
        if (a->uc[i]) {
            d->MemUnmap((CUdeviceptr)(uintptr_t)a->uc[i], a->bytes);
            d->MemAddressFree((CUdeviceptr)(uintptr_t)a->uc[i], a->bytes);
        }
    }
    if (a->mch) d->MemRelease((CUmemGenericAllocationHandle)a->mch);
    for (int i = 0; i < a->n; ++i)
        if (a->h[i]) d->MemRelease((CUmemGenericAllocationHandle)a->h[i]);
    *a = SymmAlloc();
}

int symm_create(const int* devices, int n, size_t bytes, bool want_mc, SymmAlloc* out) {
    *out = SymmAlloc();
    const drv::Api* d = drv::api();
    if (!d) {
        set_error("symmetric memory: CUDA driver entry points are not available");
        return BQ_ERR_UNSUPPORTED;
    }
    if (n < 1 || n > BQ_MAX_WORLD || bytes == 0) return BQ_ERR_INVALID;
    const bool mc_ok = want_mc && n > 1 && symm_multicast_supported(devices, n);

    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    size_t gran = 0;
    for (int i = 0; i < n; ++i) {
        prop.location.id = devices[i];
        size_t g = 0;
        BQ_DRV(d->MemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
        if (g > gran) gran = g;
    }
    CUmulticastObjectProp mcp;
    memset(&mcp, 0, sizeof(mcp));
    mcp.numDevices = unsigned(n);
    mcp.handleTypes = 0;
    mcp.flags = 0;
    if (mc_ok) {
        mcp.size = round_up(bytes, gran);
        size_t g = 0;
        if (d->MulticastGetGranularity(&g, &mcp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && g > gran) gran = g;
    }
    const size_t size = round_up(bytes, gran);
    out->n = n;
    out->bytes = size;
    for (int i = 0; i < n; ++i) out->devices[i] = devices[i];

    for (int i = 0; i < n; ++i) {
        prop.location.id = devices[i];
        CUmemGenericAllocationHandle h = 0;
        const CUresult r = d->MemCreate(&h, size, &prop, 0);
        if (r != CUDA_SUCCESS) {
            set_error("cuMemCreate(device %d, %zu bytes) failed: %s", devices[i], size, drv::error_string(r));
            symm_destroy(out);
            return BQ_ERR_CUDA;
        }
        out->h[i] = (unsigned long long)h;
        const int rc = map_for_all(d, h, size, gran, devices, n, &out->uc[i]);
        if (rc != BQ_OK) {
            symm_destroy(out);
            return rc;
        }
        int prev = -1;
        cudaGetDevice(&prev);
        cudaSetDevice(devices[i]);
        const cudaError_t e = cudaMemset(out->uc[i], 0, size);
        cudaDeviceSynchronize();
        if (prev >= 0) cudaSetDevice(prev);
        if (e != cudaSuccess) {
            set_error("clearing symmetric memory failed: %s", cudaGetErrorString(e));
            symm_destroy(out);
            return BQ_ERR_CUDA;
        }
    }
    if (mc_ok) {
        // The multicast part is best effort: a fabric that refuses it leaves plain symmetric memory.
        mcp.size = size;
        CUmemGenericAllocationHandle mch = 0;
        bool ok = d->MulticastCreate(&mch, &mcp) == CUDA_SUCCESS;
        if (ok) out->mch = (unsigned long long)mch;
        for (int i = 0; ok && i < n; ++i) {
            CUdevice dev;
            ok = d->DeviceGet(&dev, devices[i]) == CUDA_SUCCESS && d->MulticastAddDevice(mch, dev) == CUDA_SUCCESS;
        }
        for (int i = 0; ok && i < n; ++i)
            ok = d->MulticastBindMem(mch, 0, (CUmemGenericAllocationHandle)out->h[i], 0, size, 0) == CUDA_SUCCESS;
        if (ok) ok = map_for_all(d, mch, size, gran, devices, n, &out->mc) == BQ_OK;
        if (!ok) {
            out->mc = nullptr;
            if (out->mch) d->MemRelease((CUmemGenericAllocationHandle)out->mch);
            out->mch = 0;
        }
    }
    return BQ_OK;
}

}  // namespace bq
