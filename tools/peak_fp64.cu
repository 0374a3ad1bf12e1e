// Measures the CUDA-core fp64 / fp32 FMA peaks of the GPU (the roofline
// denominators MEASURED_PEAKS.json does not carry): independent FMA chains,
// 8 per thread, 1024 threads x 2 CTAs per SM, timed with CUDA events.
#include <cuda_runtime.h>
#include <stdio.h>

template <class T>
__global__ void __launch_bounds__(512) fma_chain(T* out, T a, T b, int iters) {
    T x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6,
      x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x0 = x0 * a + b; x1 = x1 * a + b; x2 = x2 * a + b; x3 = x3 * a + b;
            x4 = x4 * a + b; x5 = x5 * a + b; x6 = x6 * a + b; x7 = x7 * a + b;
        }
    }
    T s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == T(-12345)) out[0] = s;
}

template <class T>
double run(const char* name, int sms) {
    T* d;
    cudaMalloc(&d, 64);
    const int iters = 4096, grid = sms * 4, block = 512;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0);
        fma_chain<T><<<grid, block>>>(d, T(1.0000001), T(1e-9), iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        double fma = double(grid) * block * iters * 64.0;
        double tf = 2.0 * fma / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    printf("{\"peak\": \"%s\", \"tflops\": %.2f, \"gfma_per_s\": %.1f}\n", name, best, best * 500.0);
    cudaFree(d);
    return best;
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", p.name, p.multiProcessorCount, p.clockRate);
    run<double>("fp64_fma", p.multiProcessorCount);
    run<float>("fp32_fma", p.multiProcessorCount);
    return 0;
}
