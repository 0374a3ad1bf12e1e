// CPU-only build: the CUDA entry points of include/b200quad.h as stubs, so that programs written
// against the full C ABI (the CLIs, user code) link and run against libb200quad_cpu.so on a box
// without a GPU or without libcudart.  bq_cuda_device_count() answers 0, which is what callers
// branch on; everything else reports BQ_ERR_UNSUPPORTED with a message.  Never part of the CUDA
// library (libb200quad.so), where ctx.cu / scan_host.cu / noalloc.cu define the real ones.
#include <stddef.h>

#include "host_common.h"

namespace {
int unsupported(const char* what) {
    bq::set_error("%s: this is the CPU-only build of b200quad (no CUDA runtime)", what);
    return BQ_ERR_UNSUPPORTED;
}
}  // namespace

extern "C" {

int bq_cuda_device_count(void) { return 0; }
int bq_ctx_create(int, bq_ctx** out) {
    if (out) *out = nullptr;
    return unsupported("bq_ctx_create");
}
void bq_ctx_destroy(bq_ctx*) {}
int bq_ctx_device(const bq_ctx*) { return -1; }
void* bq_ctx_stream(const bq_ctx*) { return nullptr; }
int bq_ctx_set_profile(bq_ctx*, const double*, int) { return unsupported("bq_ctx_set_profile"); }
int bq_ctx_configure(bq_ctx*, int, int, int) { return unsupported("bq_ctx_configure"); }
int bq_ctx_set_timeout_ms(bq_ctx*, int) { return unsupported("bq_ctx_set_timeout_ms"); }
int bq_integrate(bq_ctx*, const bq_job*, bq_result*) { return unsupported("bq_integrate"); }
int bq_integrate_value(bq_ctx*, const bq_job*, double*, bq_result*) { return unsupported("bq_integrate_value"); }
int bq_integrate_async(bq_ctx*, const bq_job*, void*) { return unsupported("bq_integrate_async"); }
int bq_result_fetch(bq_ctx*, bq_result*) { return unsupported("bq_result_fetch"); }
int bq_integrate_device_async(bq_ctx*, const bq_job*, double*, void*) { return unsupported("bq_integrate_device_async"); }
int bq_device_result_fetch(bq_ctx*, bq_result*) { return unsupported("bq_device_result_fetch"); }
const bq_result* bq_ctx_device_result(const bq_ctx*) { return nullptr; }
int bq_cumulative(bq_ctx*, const bq_scan_job*, double*, double*, double*, bq_scan_result*) { return unsupported("bq_cumulative"); }
int bq_cumulative_async(bq_ctx*, const bq_scan_job*, double*, double*, double*, void*) { return unsupported("bq_cumulative_async"); }
int bq_scan_result_fetch(bq_ctx*, bq_scan_result*) { return unsupported("bq_scan_result_fetch"); }
int bq_ctx_reserve_scan(bq_ctx*, uint64_t) { return unsupported("bq_ctx_reserve_scan"); }
uint64_t bq_noalloc_max_samples(void) { return BQ_NOALLOC_MAX_SAMPLES; }
int bq_cumulative_noalloc(int, const bq_scan_job*, double*, double*, bq_scan_result*, void*) { return unsupported("bq_cumulative_noalloc"); }
int bq_ctx_export_comm(bq_ctx*, unsigned char*) { return unsupported("bq_ctx_export_comm"); }
int bq_ctx_attach_peers(bq_ctx*, int, int, const unsigned char*) { return unsupported("bq_ctx_attach_peers"); }
int bq_ctx_reset_comm(bq_ctx*) { return unsupported("bq_ctx_reset_comm"); }
int bq_ctx_adopt_comm(bq_ctx*, int, int, void* const*, void*, size_t) { return unsupported("bq_ctx_adopt_comm"); }
int bq_ctx_set_nvls(bq_ctx*, int) { return unsupported("bq_ctx_set_nvls"); }
int bq_ctx_set_scan_multicast(bq_ctx*, double*, double*) { return unsupported("bq_ctx_set_scan_multicast"); }
int bq_ipc_export(void*, unsigned char*) { return unsupported("bq_ipc_export"); }
int bq_ipc_export_at(void*, unsigned char*, uint64_t*) { return unsupported("bq_ipc_export_at"); }
int bq_ipc_open(const unsigned char*, void**) { return unsupported("bq_ipc_open"); }
int bq_ipc_close(void*) { return unsupported("bq_ipc_close"); }
int bq_ctx_set_scan_peers(bq_ctx*, double* const*, double* const*) { return unsupported("bq_ctx_set_scan_peers"); }
int bq_world_create(int, const int*, bq_world** out) {
    if (out) *out = nullptr;
    return unsupported("bq_world_create");
}
void bq_world_destroy(bq_world*) {}
int bq_world_size(const bq_world*) { return 0; }
bq_ctx* bq_world_ctx(bq_world*, int) { return nullptr; }
bq_ctx* bq_default_ctx(int) { unsupported("bq_default_ctx"); return nullptr; }
bq_world* bq_default_world(int) { unsupported("bq_default_world"); return nullptr; }
int bq_set_dist_ctx(bq_ctx*) { return unsupported("bq_set_dist_ctx"); }
bq_ctx* bq_dist_ctx(void) { return nullptr; }
int bq_ctx_world(const bq_ctx*) { return 0; }
int bq_ctx_rank(const bq_ctx*) { return -1; }
int bq_world_has_multicast(const bq_world*) { return 0; }
int bq_world_set_nvls(bq_world*, int) { return unsupported("bq_world_set_nvls"); }
int bq_world_alloc_tables(bq_world*, uint64_t, double**, double**) { return unsupported("bq_world_alloc_tables"); }
void bq_world_free_tables(bq_world*) {}
double* bq_world_table_multicast(bq_world*, const double*) { return nullptr; }
int bq_world_integrate(bq_world*, const bq_job*, bq_result*) { return unsupported("bq_world_integrate"); }
int bq_world_cumulative(bq_world*, const bq_scan_job*, double* const*, double* const*, bq_scan_result*) { return unsupported("bq_world_cumulative"); }
int bq_integrate_local_async(bq_ctx*, const bq_job*, int, int, double*, void*) { return unsupported("bq_integrate_local_async"); }
int bq_device_alloc(int, size_t, void** out) {
    if (out) *out = nullptr;
    return unsupported("bq_device_alloc");
}
int bq_device_free(int, void*) { return unsupported("bq_device_free"); }
int bq_device_read(int, void*, const void*, size_t) { return unsupported("bq_device_read"); }
int bq_ctx_trace(bq_ctx*, int) { return unsupported("bq_ctx_trace"); }
int bq_ctx_trace_read(bq_ctx*, uint64_t*, int) { return 0; }
int bq_ctx_inject_fault(bq_ctx*, int) { return unsupported("bq_ctx_inject_fault"); }

}  // extern "C"
