// TORCH_LIBRARY(b200quad): the thin PyTorch layer over the C ABI.
//
//   b200quad::integrate(int integrand, float a, float b, int n, int rule, int dtype,
//                       int eval, int device) -> Tensor
//       0-dim float64 tensor on `device` (CUDA index, or -1 for the CPU backend).
//       Enqueued on torch's current CUDA stream: ONE fused kernel (the job rides in
//       its launch parameters) that stores the integral into the output tensor.
//   b200quad::cumulative(int integrand, float x0, float dx, int n, int device) -> (Tensor P, Tensor Q)
//
// The reference has no library surface at all (three main()s); this is the
// "torch op" BASELINE.json asks for.
#include <ATen/ATen.h>
#include <ATen/Parallel.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <cuda_runtime_api.h>
#include <torch/library.h>

#include <mutex>
#include <unordered_map>

#include "b200quad.h"

namespace {

std::mutex g_mu;
std::unordered_map<int, bq_ctx*> g_ctx;

bq_ctx* ctx_for(int device) {
    std::lock_guard<std::mutex> lock(g_mu);
    auto it = g_ctx.find(device);
    if (it != g_ctx.end()) return it->second;
    bq_ctx* c = nullptr;
    const int rc = bq_ctx_create(device, &c);
    TORCH_CHECK(rc == BQ_OK, "b200quad: bq_ctx_create(", device, ") failed: ", bq_last_error());
    g_ctx[device] = c;
    return c;
}

bq_job make_job(int64_t integrand, double a, double b, int64_t n, int64_t rule, int64_t dtype,
                int64_t eval) {
    bq_job j{};
    j.integrand = int32_t(integrand);
    j.rule = int32_t(rule);
    j.dtype = int32_t(dtype);
    j.eval = int32_t(eval);
    j.a = a;
    j.b = b;
    j.n = uint64_t(n);
    return j;
}

at::Tensor integrate_op(int64_t integrand, double a, double b, int64_t n, int64_t rule,
                        int64_t dtype, int64_t eval, int64_t device) {
    const bq_job job = make_job(integrand, a, b, n, rule, dtype, eval);
    if (device < 0) {
        bq_result r{};
        const int rc = bq_cpu_integrate(&job, int(at::get_num_threads()), &r);
        TORCH_CHECK(rc == BQ_OK, "b200quad: ", bq_last_error());
        return at::scalar_tensor(r.value, at::dtype(at::kDouble));
    }
    c10::cuda::CUDAGuard guard(static_cast<c10::DeviceIndex>(device));
    bq_ctx* c = ctx_for(int(device));
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(static_cast<c10::DeviceIndex>(device));
    // ONE kernel: it stores the integral straight into the output tensor.
    at::Tensor out = at::empty({}, at::dtype(at::kDouble).device(at::kCUDA, device));
    const int rc = bq_integrate_device_async(c, &job, out.data_ptr<double>(), stream);
    TORCH_CHECK(rc == BQ_OK, "b200quad: ", bq_last_error());
    return out;
}

std::tuple<at::Tensor, at::Tensor> cumulative_op(int64_t integrand, double x0, double dx, int64_t n,
                                                 int64_t device) {
    bq_scan_job job{};
    job.integrand = int32_t(integrand);
    job.x0 = x0;
    job.dx = dx;
    job.n = uint64_t(n);
    if (device < 0) {
        at::Tensor P = at::empty({n}, at::dtype(at::kDouble));
        at::Tensor Q = at::empty({n}, at::dtype(at::kDouble));
        bq_scan_result r{};
        const int rc = bq_cpu_cumulative(&job, int(at::get_num_threads()), P.data_ptr<double>(),
                                         Q.data_ptr<double>(), nullptr, &r);
        TORCH_CHECK(rc == BQ_OK, "b200quad: ", bq_last_error());
        return {P, Q};
    }
    c10::cuda::CUDAGuard guard(static_cast<c10::DeviceIndex>(device));
    bq_ctx* c = ctx_for(int(device));
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(static_cast<c10::DeviceIndex>(device));
    at::Tensor P = at::empty({n}, at::dtype(at::kDouble).device(at::kCUDA, device));
    at::Tensor Q = at::empty({n}, at::dtype(at::kDouble).device(at::kCUDA, device));
    const int rc = bq_cumulative_async(c, &job, P.data_ptr<double>(), Q.data_ptr<double>(), nullptr,
                                       stream);
    TORCH_CHECK(rc == BQ_OK, "b200quad: ", bq_last_error());
    return {P, Q};
}

}  // namespace

TORCH_LIBRARY(b200quad, m) {
    m.def("integrate(int integrand, float a, float b, int n, int rule=0, int dtype=0, int eval=0, "
          "int device=0) -> Tensor");
    m.def("cumulative(int integrand, float x0, float dx, int n, int device=0) -> (Tensor, Tensor)");
}

TORCH_LIBRARY_IMPL(b200quad, CompositeExplicitAutograd, m) {
    m.impl("integrate", integrate_op);
    m.impl("cumulative", cumulative_op);
}
