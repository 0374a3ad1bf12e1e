// What limits the fp32 ROTATE loop (acc_j = fma(C, Q_j, fma(S, P_j, acc_j)), 8 columns per thread, row
// values (S, C) from shared memory, the fp32 runs promoted to fp64 every kPromoteRows rows)?  Variants of
// the same arithmetic, as a fraction of the fp32 FMA peak (35 381 G FMA lanes/s, tools/peak_fp64.cu).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o bin/ffma_bench tools/ffma_bench.cu
#include <cuda_runtime.h>
#include <stdio.h>

__device__ __forceinline__ unsigned long long pack(float a, float b) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ void unpack(unsigned long long v, float& a, float& b) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ void fma2(unsigned long long& acc, unsigned long long x, unsigned long long y) {
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(x), "l"(y));
}
__device__ __forceinline__ unsigned long long add2(unsigned long long x, unsigned long long y) {
    unsigned long long r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
    return r;
}
__device__ __forceinline__ unsigned long long sub2(unsigned long long x, unsigned long long y) {
    unsigned long long r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
    return r;
}

// MODE 0: scalar FFMA, no promotion             MODE 1: packed FFMA2, no promotion
// MODE 2: FFMA2, promotion to fp64 every PROMO rows (the kernel's loop)
// MODE 3: FFMA2, every PROMO rows the run goes into a packed fp32 (hi, lo) pair by a branch-free
//         TwoSum (6 packed adds per column pair); promotion of the pair to fp64 every 16 * PROMO rows
// MODE 6: as 2 with the two products of a sample in separate accumulators (independent FMA chains) that
//         meet in one packed add at the promotion (the kernel's loop since the error-diffusion change)
// MODE 4: FFMA2 peak probe: 16 independent packed accumulators, both multiplicands the same register
// MODE 5: scalar FFMA peak probe: 16 independent accumulators, multiplicands the same register
template <int MODE, int PROMO>
__global__ void __launch_bounds__(256, 2) k(double* out, int rows, float h) {
    __shared__ float2 sSC[130];
    float P[8], Q[8], a[8];
    double d[8];
    unsigned long long P2[4], Q2[4], a2[4], hi2[4], lo2[4], x2[16];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        __sincosf(h * (threadIdx.x + 256 * j), &Q[j], &P[j]);
        a[j] = 0.f;
        d[j] = 0.0;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        P2[j] = pack(P[2 * j], P[2 * j + 1]);
        Q2[j] = pack(Q[2 * j], Q[2 * j + 1]);
        a2[j] = pack(0.f, 0.f);
        hi2[j] = pack(0.f, 0.f);
        lo2[j] = pack(0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) x2[j] = pack(0.f, 1e-30f * j);
    if (threadIdx.x < 130) sSC[threadIdx.x] = make_float2(__sinf(0.001f * threadIdx.x + blockIdx.x), __cosf(0.001f * threadIdx.x));
    __syncthreads();
    float2 cur = sSC[0];
    for (int r0 = 0; r0 < rows; r0 += 128) {
#pragma unroll 16
        for (int r = 0; r < 128; ++r) {
            const float2 nxt = sSC[r + 1];
            if (MODE == 0) {
#pragma unroll
                for (int j = 0; j < 8; ++j) a[j] = fmaf(cur.x, P[j], a[j]);
#pragma unroll
                for (int j = 0; j < 8; ++j) a[j] = fmaf(cur.y, Q[j], a[j]);
            } else if (MODE == 4) {
                const unsigned long long s2 = pack(cur.x, cur.x);
#pragma unroll
                for (int j = 0; j < 8; ++j) fma2(x2[j], s2, s2);
            } else if (MODE == 5) {
                float* xf = reinterpret_cast<float*>(x2);
#pragma unroll
                for (int j = 0; j < 16; ++j) xf[j] = fmaf(cur.x, cur.x, xf[j]);
            } else if (MODE == 6) {
                const unsigned long long s2 = pack(cur.x, cur.x), c2 = pack(cur.y, cur.y);
#pragma unroll
                for (int j = 0; j < 4; ++j) fma2(a2[j], s2, P2[j]);
#pragma unroll
                for (int j = 0; j < 4; ++j) fma2(hi2[j], c2, Q2[j]);
                if ((r % PROMO) == PROMO - 1) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float lo, hi;
                        unpack(add2(a2[j], hi2[j]), lo, hi);
                        d[2 * j] += double(lo);
                        d[2 * j + 1] += double(hi);
                        a2[j] = pack(0.f, 0.f);
                        hi2[j] = pack(0.f, 0.f);
                    }
                }
            } else {
                const unsigned long long s2 = pack(cur.x, cur.x), c2 = pack(cur.y, cur.y);
#pragma unroll
                for (int j = 0; j < 4; ++j) fma2(a2[j], s2, P2[j]);
#pragma unroll
                for (int j = 0; j < 4; ++j) fma2(a2[j], c2, Q2[j]);
                if ((MODE == 2 || MODE == 3) && (r % PROMO) == PROMO - 1) {
                    if (MODE == 2) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            float lo, hi;
                            unpack(a2[j], lo, hi);
                            d[2 * j] += double(lo);
                            d[2 * j + 1] += double(hi);
                            a2[j] = pack(0.f, 0.f);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {   // TwoSum(hi, run): s = hi + x; bb = s - hi; err = (hi - (s - bb)) + (x - bb)
                            const unsigned long long s = add2(hi2[j], a2[j]);
                            const unsigned long long bb = sub2(s, hi2[j]);
                            const unsigned long long e = add2(sub2(hi2[j], sub2(s, bb)), sub2(a2[j], bb));
                            hi2[j] = s;
                            lo2[j] = add2(lo2[j], e);
                            a2[j] = pack(0.f, 0.f);
                        }
                        if ((r % (16 * PROMO)) == 16 * PROMO - 1 || r == 127) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                float l0, l1, h0, h1;
                                unpack(hi2[j], h0, h1);
                                unpack(lo2[j], l0, l1);
                                d[2 * j] += double(h0) + double(l0);
                                d[2 * j + 1] += double(h1) + double(l1);
                                hi2[j] = pack(0.f, 0.f);
                                lo2[j] = pack(0.f, 0.f);
                            }
                        }
                    }
                }
            }
            cur = nxt;
        }
    }
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += d[j] + double(a[j]);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float u, v;
        unpack(a2[j], u, v);
        s += double(u) + double(v);
        unpack(hi2[j], u, v);
        s += double(u) + double(v);
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        float u, v;
        unpack(x2[j], u, v);
        s += double(u) + double(v);
    }
    if (s == -1.2345) out[0] = s;
}

// MODE 7 (own kernel): the row constants ride in the kernel parameters and enter the packed FMAs as
// uniform operands; the per-column pairs (A_j, A_j+1), (B_j, B_j+1) sit in registers (the fp64 loop's form).
struct RowParamsF {
    double* out;
    int groups, nt;
    float rc[128][2];
};
template <int PROMO>
__global__ void __launch_bounds__(256, 2) k_uniform(const __grid_constant__ RowParamsF p) {
    unsigned long long A2[4], B2[4], a2[4], b2[4];
    double d[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        A2[j] = pack(1.0f + 1e-3f * threadIdx.x, 0.5f + j);
        B2[j] = pack(1e-3f * j, 2e-3f * threadIdx.x);
        a2[j] = pack(0.f, 0.f);
        b2[j] = pack(0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) d[j] = 0.0;
    for (int g = 0; g < p.groups; ++g) {
#pragma unroll 1
        for (int kk = 0; kk < p.nt; ++kk) {
#pragma unroll
            for (int r = 0; r < 32; ++r) {
                const float c = p.rc[kk * 32 + r][0], s = p.rc[kk * 32 + r][1];
                const unsigned long long c2 = pack(c, c), s2 = pack(s, s);
#pragma unroll
                for (int j = 0; j < 4; ++j) fma2(a2[j], c2, A2[j]);
#pragma unroll
                for (int j = 0; j < 4; ++j) fma2(b2[j], s2, B2[j]);
                if ((r % PROMO) == PROMO - 1) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float lo, hi;
                        unpack(a2[j], lo, hi);
                        d[2 * j] += double(lo);
                        d[2 * j + 1] += double(hi);
                        a2[j] = pack(0.f, 0.f);
                    }
                }
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float lo, hi;
            unpack(b2[j], lo, hi);
            d[2 * j] += double(lo);
            d[2 * j + 1] += double(hi);
            b2[j] = pack(0.f, 0.f);
        }
    }
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += d[j];
    if (s == -1.2345) p.out[0] = s;
}

template <int PROMO>
void run_uniform(int sms) {
    RowParamsF p;
    cudaMalloc(&p.out, 64);
    p.groups = (1 << 15) / 128;
    p.nt = 4;
    for (int i = 0; i < 128; ++i) p.rc[i][0] = 1.0f - 1e-6f * i, p.rc[i][1] = 1e-3f * i;
    for (int ctas = 1; ctas <= 2; ++ctas) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        float best = 1e9;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            k_uniform<PROMO><<<sms * ctas, 256>>>(p);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        const double fmas = double(sms) * ctas * 256.0 * (1 << 15) * 16.0;
        printf("{\"mode\": 7, \"what\": \"row constants in kernel parameters (uniform operands), split accumulators\", \"promote_rows\": %d, \"ctas_per_sm\": %d, \"ms\": %.4f, \"frac_of_fp32_fma_peak\": %.3f}\n",
               PROMO, ctas, best, fmas / (best * 1e-3) / 1e9 / 35381.0);
    }
    cudaFree(p.out);
}

template <int MODE, int PROMO>
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
            k<MODE, PROMO><<<sms * ctas, 256>>>(d, rows, 1e-6f);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        const double fmas = double(sms) * ctas * 256.0 * rows * 16.0;   // every mode: 16 fp32 FMAs per thread per row
        printf("{\"mode\": %d, \"what\": \"%s\", \"promote_rows\": %d, \"ctas_per_sm\": %d, \"ms\": %.4f, \"frac_of_fp32_fma_peak\": %.3f}\n",
               MODE, what, PROMO, ctas, best, fmas / (best * 1e-3) / 1e9 / 35381.0);
    }
    cudaFree(d);
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    run<5, 1>(sms, "scalar FFMA probe: 16 independent accumulators, one multiplicand register");
    run<4, 1>(sms, "FFMA2 probe: 8 independent packed accumulators, one multiplicand register");
    run<0, 1>(sms, "scalar FFMA row loop, no promotion");
    run<1, 1>(sms, "FFMA2 row loop, no promotion");
    run<2, 16>(sms, "FFMA2 row loop, fp64 promotion every 16 rows (the kernel)");
    run<2, 32>(sms, "FFMA2 row loop, fp64 promotion every 32 rows");
    run<2, 8>(sms, "FFMA2 row loop, fp64 promotion every 8 rows");
    run<6, 16>(sms, "FFMA2 row loop, split accumulators, fp64 promotion every 16 rows");
    run<6, 32>(sms, "FFMA2 row loop, split accumulators, fp64 promotion every 32 rows");
    run_uniform<16>(sms);
    run_uniform<32>(sms);
    run<3, 16>(sms, "FFMA2 row loop, packed fp32 TwoSum every 16 rows, fp64 every 128");
    run<3, 8>(sms, "FFMA2 row loop, packed fp32 TwoSum every 8 rows, fp64 every 128");
    return 0;
}
