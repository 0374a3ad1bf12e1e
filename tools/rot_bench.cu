// Inner-loop study for the ROTATE path: how close can DMUL+DFMA+DADD per point
// get to the fp64 issue peak, as a function of columns per thread, unroll,
// software prefetch of the row base, the per-tile barrier and CTAs per SM.
#include <cuda_runtime.h>
#include <stdio.h>

template <int COLS, int UNROLL, bool PREFETCH, bool SYNC, int ROWS = 64>
__global__ void __launch_bounds__(256) rot(double* out, int tiles, double h) {
    __shared__ double2 sSC[2][64];
    double cj[COLS], sj[COLS], acc[COLS];
#pragma unroll
    for (int j = 0; j < COLS; ++j) {
        sincos(h * (threadIdx.x + 256 * j), &sj[j], &cj[j]);
        acc[j] = 0.0;
    }
    if (threadIdx.x < 64) {
        double s, c;
        sincos(0.001 * threadIdx.x + blockIdx.x, &s, &c);
        sSC[0][threadIdx.x] = make_double2(s, c);
        sSC[1][threadIdx.x] = make_double2(c, s);
    }
    __syncthreads();
    double total = 0.0;
    for (int t = 0; t < tiles; ++t) {
        const double2* base = sSC[t & 1];
        if (PREFETCH) {
            double2 cur = base[0];
#pragma unroll UNROLL
            for (int r = 0; r < ROWS; ++r) {
                const double2 nxt = base[(r + 1) & 63];
#pragma unroll
                for (int j = 0; j < COLS; ++j) acc[j] += fma(cur.y, sj[j], cur.x * cj[j]);
                cur = nxt;
            }
        } else {
#pragma unroll UNROLL
            for (int r = 0; r < ROWS; ++r) {
                const double2 sc = base[r];
#pragma unroll
                for (int j = 0; j < COLS; ++j) acc[j] += fma(sc.y, sj[j], sc.x * cj[j]);
            }
        }
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < COLS; ++j) s += acc[j], acc[j] = 0.0;
        total += s;
        if (SYNC) __syncthreads();
    }
    if (total == -1.2345) out[0] = total;
}

template <int COLS, int UNROLL, bool PREFETCH, bool SYNC, int ROWS = 64>
void run(int sms, double peak_gdp) {
    double* d;
    cudaMalloc(&d, 64);
    for (int ctas = 1; ctas <= 4; ++ctas) {
        int maxb = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&maxb, rot<COLS, UNROLL, PREFETCH, SYNC, ROWS>, 256, 0);
        if (ctas > maxb) break;
        const int tiles = 2048 / COLS * 4;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        float best = 1e9;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            rot<COLS, UNROLL, PREFETCH, SYNC, ROWS><<<sms * ctas, 256>>>(d, tiles, 1e-9);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        const double pts = double(sms) * ctas * 256.0 * tiles * double(ROWS) * COLS;
        const double gps = pts / (best * 1e-3) / 1e9;
        printf("{\"rows\": %d, \"cols\": %d, \"unroll\": %d, \"prefetch\": %d, \"sync\": %d, \"ctas_per_sm\": %d, "
               "\"gpts_per_s\": %.1f, \"frac_of_dp_issue_peak\": %.3f}\n",
               ROWS, COLS, UNROLL, int(PREFETCH), int(SYNC), ctas, gps, gps * 3.0 / peak_gdp);
    }
    cudaFree(d);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    const double peak = 18563.0;   // G DP-instr-lanes/s measured by tools/peak_fp64.cu
    run<8, 4, true, true, 64>(sms, peak);
    run<8, 4, true, true, 32>(sms, peak);
    run<8, 4, true, true, 16>(sms, peak);
    run<8, 4, true, false, 32>(sms, peak);
    run<8, 4, false, true, 32>(sms, peak);
    run<8, 8, true, true, 32>(sms, peak);
    return 0;
}
