// What limits the 2-DFMA ROTATE loop (acc_j = fma(C, Q_j, fma(S, P_j, acc_j)), 8 columns, row values
// (S, C) from shared memory)?  Variants of the same arithmetic with different operand patterns,
// as a fraction of the fp64 issue peak (18 563 G instruction-lanes/s measured by tools/peak_fp64.cu).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o bin/dfma_bench tools/dfma_bench.cu
#include <cuda_runtime.h>
#include <stdio.h>

// MODE 0: the kernel's loop: per row one LDS.128 (S, C), 16 DFMAs, software-prefetched
// MODE 1: no shared memory: S, C advance in registers (two extra DADDs per row)
// MODE 2: as 0 but the two DFMAs of a column are issued back to back (S then C per column)
// MODE 3: as 0 with the product operands swapped (fma(P_j, S, acc) instead of fma(S, P_j, acc))
// MODE 4: as 0 with 16 accumulators (two rows in flight: even rows into acc[0..7], odd into acc[8..15])
// MODE 5: row values from __constant__ memory at a warp-uniform address (a probe: can a uniform / constant
//         operand, which takes one register read off each DFMA, lift the ceiling?)
__constant__ double2 cSC[130];

template <int MODE, int UNROLL>
__global__ void __launch_bounds__(256, 2) k(double* out, int rows, double h) {
    __shared__ double2 sSC[130];
    double P[8], Q[8], acc[16];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        sincos(h * (threadIdx.x + 256 * j), &Q[j], &P[j]);
        acc[j] = 0.0;
        acc[j + 8] = 0.0;
    }
    if (threadIdx.x < 130) sSC[threadIdx.x] = make_double2(sin(0.001 * threadIdx.x + blockIdx.x), cos(0.001 * threadIdx.x));
    __syncthreads();
    const double2* base = sSC;
    double2 cur = base[0];
    for (int r0 = 0; r0 < rows; r0 += 128) {
#pragma unroll UNROLL
        for (int r = 0; r < 128; ++r) {
            double2 nxt;
            if (MODE == 1) {
                nxt = make_double2(cur.x + h, cur.y - h);
            } else if (MODE == 5) {
                nxt = cSC[r + 1];
            } else {
                nxt = base[r + 1];
            }
            if (MODE == 2) {
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fma(cur.y, Q[j], fma(cur.x, P[j], acc[j]));
            } else if (MODE == 3) {
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fma(P[j], cur.x, acc[j]);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fma(Q[j], cur.y, acc[j]);
            } else if (MODE == 4) {
                const int o = (r & 1) * 8;
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j + o] = fma(cur.x, P[j], acc[j + o]);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j + o] = fma(cur.y, Q[j], acc[j + o]);
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fma(cur.x, P[j], acc[j]);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fma(cur.y, Q[j], acc[j]);
            }
            cur = nxt;
        }
    }
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += acc[j];
    if (s == -1.2345) out[0] = s;
}

// MODE 6 (own kernel): the row constants ride in the kernel parameters (constant bank 0), the per-column
// values A_j, B_j sit in registers and are re-derived once per group of 4 x 32 rows — the form the library's
// fp64 loop takes: acc_j = fma(B_j, s_r, fma(A_j, c_r, acc_j)) with c_r, s_r as uniform-register operands.
struct RowParams {
    double* out;
    int groups, nt;
    double rc[128][2];
};
__global__ void __launch_bounds__(256, 2) k_uniform(const __grid_constant__ RowParams p) {
    double P[8], Q[8], A[8], B[8], acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        sincos(1e-9 * (threadIdx.x + 256 * j), &Q[j], &P[j]);
        acc[j] = 0.0;
    }
    double sb = 0.3 + 1e-3 * blockIdx.x, cb = 0.9;
    for (int g = 0; g < p.groups; ++g) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {      // 4 fp64 instructions per column per group
            A[j] = fma(sb, P[j], cb * Q[j]);
            B[j] = fma(cb, P[j], -(sb * Q[j]));
        }
#pragma unroll 1
        for (int kk = 0; kk < p.nt; ++kk) {
#pragma unroll
            for (int r = 0; r < 32; ++r) {
                const double c = p.rc[kk * 32 + r][0], s = p.rc[kk * 32 + r][1];
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fma(A[j], c, acc[j]);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fma(B[j], s, acc[j]);
            }
        }
        sb += 1e-7, cb -= 1e-7;
    }
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += acc[j];
    if (s == -1.2345) p.out[0] = s;
}

void run_uniform(int sms) {
    RowParams p;
    cudaMalloc(&p.out, 64);
    p.groups = (1 << 15) / 128;
    p.nt = 4;
    for (int i = 0; i < 128; ++i) p.rc[i][0] = 1.0 - 1e-6 * i, p.rc[i][1] = 1e-3 * i;
    for (int ctas = 1; ctas <= 2; ++ctas) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        float best = 1e9;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            k_uniform<<<sms * ctas, 256>>>(p);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        const double dfma = double(sms) * ctas * 256.0 * (1 << 15) * 16.0;
        printf("{\"mode\": 6, \"what\": \"row constants in kernel parameters (uniform operands), A/B per column re-derived per 128 rows\", \"unroll\": 32, \"ctas_per_sm\": %d, \"ms\": %.4f, \"frac_of_fp64_issue_peak\": %.3f}\n",
               ctas, best, dfma / (best * 1e-3) / 1e9 / 18563.0);
    }
    cudaFree(p.out);
}

template <int MODE, int UNROLL>
void run(int sms, const char* what) {
    double* d;
    cudaMalloc(&d, 64);
    for (int ctas = 1; ctas <= 2; ++ctas) {
        const int rows = 1 << 15;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        float best = 1e9;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            k<MODE, UNROLL><<<sms * ctas, 256>>>(d, rows, 1e-9);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        const double dfma = double(sms) * ctas * 256.0 * rows * 16.0;
        printf("{\"mode\": %d, \"what\": \"%s\", \"unroll\": %d, \"ctas_per_sm\": %d, \"ms\": %.4f, \"frac_of_fp64_issue_peak\": %.3f}\n",
               MODE, what, UNROLL, ctas, best, dfma / (best * 1e-3) / 1e9 / 18563.0);
    }
    cudaFree(d);
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    run<0, 16>(sms, "kernel loop (LDS row values)");
    run<0, 32>(sms, "kernel loop, unroll 32");
    run<1, 16>(sms, "row values in registers (no LDS)");
    run<2, 16>(sms, "S and C of a column back to back");
    run<3, 16>(sms, "product operands swapped");
    run<4, 16>(sms, "16 accumulators");
    {
        double2 hsc[130];
        for (int i = 0; i < 130; ++i) hsc[i] = make_double2(0.001 * i, 1.0 - 0.001 * i);
        cudaMemcpyToSymbol(cSC, hsc, sizeof(hsc));
    }
    run<5, 128>(sms, "row values from constant memory (uniform operand), fully unrolled");
    run_uniform(sms);
    return 0;
}
