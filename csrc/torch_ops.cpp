// TORCH_LIBRARY(b200quad): the thin PyTorch layer over the C ABI.
//
//   b200quad::integrate(int integrand, float a, float b, int n, int rule, int dtype,
//                       int eval, int device, int world, float[] params) -> Tensor
//       0-dim float64 tensor on `device` (CUDA index, or -1 for the CPU backend).
//       Enqueued on torch's current CUDA stream: ONE fused kernel (the job rides in
//       its launch parameters) that stores the integral into the output tensor.
//       world > 1 (the rank count is an input of the reference's runs: riemann.cpp:63-65,
//       4main.c:70-71): under torch.distributed this rank's kernel of the collective call;
//       in a single process the process-wide world of that many GPUs.
//       params: integrand parameters (bq_job.p; empty = the reference's constants).
//   b200quad::cumulative(int integrand, float x0, float dx, int n, int device, int world,
//                        float[] params) -> (Tensor P, Tensor Q)
//
// Contexts are the process-wide ones (bq_default_ctx / bq_dist_ctx): the same objects the
// Python layer uses, so e.g. Context.set_profile is seen here.
//
// The reference has no library surface at all (three main()s); this is the
// "torch op" BASELINE.json asks for.
#include <ATen/ATen.h>
#include <ATen/Parallel.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <cuda_runtime_api.h>
#include <torch/library.h>

#include <algorithm>

#include "b200quad.h"

namespace {

// world <= 1: the process-wide context of the device (the one parallel.context() wraps).
bq_ctx* ctx_for(int device) {
    bq_ctx* c = bq_default_ctx(device);
    TORCH_CHECK(c != nullptr, "b200quad: no context for device ", device, ": ", bq_last_error());
    return c;
}

// world > 1 under torch.distributed: the context DistWorld registered for this rank.
bq_ctx* dist_ctx_for(int64_t world) {
    bq_ctx* c = bq_dist_ctx();
    if (c && bq_ctx_world(c) == world) return c;
    return nullptr;
}

// torch's default stream is the NULL handle, which the C ABI reads as "the context's own stream":
// name it explicitly so the work is ordered with everything else torch does on that stream.
cudaStream_t current_stream(int device) {
    cudaStream_t s = c10::cuda::getCurrentCUDAStream(static_cast<c10::DeviceIndex>(device));
    return s ? s : cudaStreamLegacy;
}

bq_job make_job(int64_t integrand, double a, double b, int64_t n, int64_t rule, int64_t dtype,
                int64_t eval, at::ArrayRef<double> params) {
    bq_job j{};
    j.integrand = int32_t(integrand);
    j.rule = int32_t(rule);
    j.dtype = int32_t(dtype);
    j.eval = int32_t(eval);
    j.a = a;
    j.b = b;
    j.n = uint64_t(n);
    TORCH_CHECK(params.size() <= 4, "b200quad: at most 4 integrand parameters");
    for (size_t i = 0; i < params.size(); ++i) j.p[i] = params[i];
    return j;
}

at::Tensor integrate_op(int64_t integrand, double a, double b, int64_t n, int64_t rule,
                        int64_t dtype, int64_t eval, int64_t device, int64_t world,
                        at::ArrayRef<double> params) {
    const bq_job job = make_job(integrand, a, b, n, rule, dtype, eval, params);
    if (device < 0) {
        bq_result r{};
        const int rc = bq_cpu_integrate(&job, int(at::get_num_threads()), &r);
        TORCH_CHECK(rc == BQ_OK, "b200quad: ", bq_last_error());
        return at::scalar_tensor(r.value, at::dtype(at::kDouble));
    }
    if (world > 1) {
        if (bq_ctx* dc = dist_ctx_for(world)) {
            // one process per GPU: this rank's kernel, enqueued on torch's current stream; the
            // in-kernel NVLink combine leaves the same bits in every rank's output tensor
            const int dev = bq_ctx_device(dc);
            c10::cuda::CUDAGuard guard(static_cast<c10::DeviceIndex>(dev));
            cudaStream_t stream = current_stream(dev);
            at::Tensor out = at::empty({}, at::dtype(at::kDouble).device(at::kCUDA, dev));
            const int rc = bq_integrate_device_async(dc, &job, out.data_ptr<double>(), stream);
            TORCH_CHECK(rc == BQ_OK, "b200quad: ", bq_last_error());
            return out;
        }
        // one process driving `world` GPUs: launch on all of them, wait, result on `device`
        bq_world* w = bq_default_world(int(world));
        TORCH_CHECK(w != nullptr, "b200quad: world of ", world, " GPUs: ", bq_last_error());
        bq_result res[BQ_MAX_WORLD];
        const int rc = bq_world_integrate(w, &job, res);
        TORCH_CHECK(rc == BQ_OK, "b200quad: ", bq_last_error());
        return at::scalar_tensor(res[0].value, at::dtype(at::kDouble).device(at::kCUDA, device));
    }
    c10::cuda::CUDAGuard guard(static_cast<c10::DeviceIndex>(device));
    bq_ctx* c = ctx_for(int(device));
    cudaStream_t stream = current_stream(int(device));
    // ONE kernel: it stores the integral straight into the output tensor.
    at::Tensor out = at::empty({}, at::dtype(at::kDouble).device(at::kCUDA, device));
    const int rc = bq_integrate_device_async(c, &job, out.data_ptr<double>(), stream);
    TORCH_CHECK(rc == BQ_OK, "b200quad: ", bq_last_error());
    return out;
}

std::tuple<at::Tensor, at::Tensor> cumulative_op(int64_t integrand, double x0, double dx, int64_t n,
                                                 int64_t device, int64_t world, at::ArrayRef<double> params) {
    bq_scan_job job{};
    job.integrand = int32_t(integrand);
    job.x0 = x0;
    job.dx = dx;
    job.n = uint64_t(n);
    TORCH_CHECK(params.size() <= 4, "b200quad: at most 4 integrand parameters");
    for (size_t i = 0; i < params.size(); ++i) job.p[i] = params[i];
    if (device < 0) {
        at::Tensor P = at::empty({n}, at::dtype(at::kDouble));
        at::Tensor Q = at::empty({n}, at::dtype(at::kDouble));
        bq_scan_result r{};
        const int rc = bq_cpu_cumulative(&job, int(at::get_num_threads()), P.data_ptr<double>(),
                                         Q.data_ptr<double>(), nullptr, &r);
        TORCH_CHECK(rc == BQ_OK, "b200quad: ", bq_last_error());
        return {P, Q};
    }
    if (world > 1) {
        // This rank's slice of the tables (replicate = 0): rank r of a world of W owns 8/W of the 8
        // fixed segments of the tile space; the carries cross NVLink inside the kernel.
        bq_ctx* dc = dist_ctx_for(world);
        TORCH_CHECK(dc != nullptr, "b200quad::cumulative(world=", world, ") needs the torch.distributed world of that "
                    "size (b200quad.parallel.dist_world()); a single process uses bq.cumulative_integrate(world=)");
        const int dev = bq_ctx_device(dc), rank = bq_ctx_rank(dc);
        const int64_t tiles = (n + 2047) / 2048, segs = 8 / world;
        const int64_t lo = (rank * segs * tiles / 8) * 2048;
        const int64_t hi = std::min<int64_t>(n, ((rank + 1) * segs * tiles / 8) * 2048);
        c10::cuda::CUDAGuard guard(static_cast<c10::DeviceIndex>(dev));
        cudaStream_t stream = current_stream(dev);
        // a rank whose segments hold no tile (tiny n) still takes part: a valid, empty slice
        const int64_t m = std::max<int64_t>(hi - lo, 0);
        // (the kernel gets the base buffers: an empty view has no address)
        at::Tensor P = at::empty({std::max<int64_t>(m, 4)}, at::dtype(at::kDouble).device(at::kCUDA, dev));
        at::Tensor Q = at::empty({std::max<int64_t>(m, 4)}, at::dtype(at::kDouble).device(at::kCUDA, dev));
        const int rc = bq_cumulative_async(dc, &job, P.data_ptr<double>(), Q.data_ptr<double>(), nullptr, stream);
        TORCH_CHECK(rc == BQ_OK, "b200quad: ", bq_last_error());
        return {P.narrow(0, 0, m), Q.narrow(0, 0, m)};
    }
    c10::cuda::CUDAGuard guard(static_cast<c10::DeviceIndex>(device));
    bq_ctx* c = ctx_for(int(device));
    cudaStream_t stream = current_stream(int(device));
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
          "int device=0, int world=1, float[] params=[]) -> Tensor");
    m.def("cumulative(int integrand, float x0, float dx, int n, int device=0, int world=1, "
          "float[] params=[]) -> (Tensor, Tensor)");
}

TORCH_LIBRARY_IMPL(b200quad, CompositeExplicitAutograd, m) {
    m.impl("integrate", integrate_op);
    m.impl("cumulative", cumulative_op);
}
