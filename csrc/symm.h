// Symmetric, multicast-capable device memory for a single-process world (one allocation per GPU,
// each mapped for every GPU, all bound to ONE NVLS multicast object).
//
// A store to the multicast address is replicated by the NVSwitch into every GPU's copy
// (multimem.st in the kernels: one instruction instead of one store per peer).  Replaces, as a
// transport, MPI_Bcast / MPI_Send of the reference (4main.c:143-157, riemann.cpp:76).
#pragma once
#include <stddef.h>

#include "../include/b200quad.h"

namespace bq {

struct SymmAlloc {
    int n = 0;
    size_t bytes = 0;                       // mapped size of each copy (granularity-rounded)
    void* uc[BQ_MAX_WORLD] = {};            // copy of rank r, addressable from every device of the world
    void* mc = nullptr;                     // multicast address (null when NVLS is unavailable)
    unsigned long long h[BQ_MAX_WORLD] = {};   // CUmemGenericAllocationHandle of each copy
    unsigned long long mch = 0;                // ... of the multicast object
    int devices[BQ_MAX_WORLD] = {};
};

// True when every listed device reports CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED.
bool symm_multicast_supported(const int* devices, int n);
// Allocates zero-filled copies on every device.  With want_mc the copies are bound to a multicast
// object; a failure of that part alone leaves out->mc null and still succeeds (plain symmetric
// memory).  Returns a bq_status; on failure nothing is left allocated.
int symm_create(const int* devices, int n, size_t bytes, bool want_mc, SymmAlloc* out);
void symm_destroy(SymmAlloc* a);

}  // namespace bq
