// Floor of what CUDA events see around ONE kernel launch that follows a memset on the same stream
// (the shape of a bench.py step): an empty kernel with the quad kernel's grid (2 CTAs per SM, cluster
// pairs, 256 threads) and with a plain grid, median of 200.  What is left of a step's event time
// after subtracting the kernel's own %globaltimer span should be close to this number.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o bin/launch_floor tools/launch_floor.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <algorithm>
#include <vector>

__global__ void empty_kernel(unsigned long long* out) {
    if (out && blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        out[0] = t;
    }
}

static float median_us(int grid, int cluster, bool flush, void* buf, size_t bytes, cudaStream_t s) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    std::vector<float> v;
    for (int i = 0; i < 210; ++i) {
        if (flush) cudaMemsetAsync(buf, 0, bytes, s);
        cudaEventRecord(e0, s);
        cudaLaunchConfig_t lc = {};
        lc.gridDim = dim3(grid);
        lc.blockDim = dim3(256);
        lc.stream = s;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = cluster;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        lc.attrs = at;
        lc.numAttrs = 1;
        cudaLaunchKernelEx(&lc, empty_kernel, (unsigned long long*)nullptr);
        cudaEventRecord(e1, s);
        cudaStreamSynchronize(s);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (i >= 10) v.push_back(ms * 1e3f);
    }
    std::sort(v.begin(), v.end());
    return v[v.size() / 2];
}

int main() {
    cudaStream_t s;
    cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    void* buf;
    const size_t bytes = 256u << 20;
    cudaMalloc(&buf, bytes);
    printf("{\"sms\": %d", sms);
    printf(", \"empty_cluster2_grid%d_after_memset_us\": %.2f", 2 * sms, median_us(2 * sms, 2, true, buf, bytes, s));
    printf(", \"empty_cluster1_grid%d_after_memset_us\": %.2f", 2 * sms, median_us(2 * sms, 1, true, buf, bytes, s));
    printf(", \"empty_cluster1_grid1_after_memset_us\": %.2f", median_us(1, 1, true, buf, bytes, s));
    printf(", \"empty_cluster2_grid%d_idle_stream_us\": %.2f", 2 * sms, median_us(2 * sms, 2, false, buf, bytes, s));
    printf(", \"empty_cluster1_grid1_idle_stream_us\": %.2f}\n", median_us(1, 1, false, buf, bytes, s));
    return 0;
}
