// Double-double ("dd") helpers shared by the CUDA kernels and the CPU backend.
//
// Every cross-thread / cross-CTA / cross-GPU combine in this library is done on
// (hi, lo) pairs with error-free transformations, so the only uncompensated
// additions are the short per-thread runs inside one tile.  This is what lets
// the library hit <=1e-15 relative error where the reference's naive fp64 sum
// is off by 1.1e-12 (riemann.cpp:31-43) to 2.5e-10 (4main.c:103-107).
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define BQ_HD __host__ __device__ __forceinline__
#else
#define BQ_HD inline
#endif

namespace bq {

struct dd {
    double hi, lo;
};

// Knuth TwoSum: s + e == a + b exactly.
BQ_HD dd two_sum(double a, double b) {
    double s = a + b;
    double bb = s - a;
    double e = (a - (s - bb)) + (b - bb);
    return dd{s, e};
}

// Dekker FastTwoSum, requires |a| >= |b| (or a == 0).
BQ_HD dd fast_two_sum(double a, double b) {
    double s = a + b;
    double e = b - (s - a);
    return dd{s, e};
}

BQ_HD dd dd_add(dd a, dd b) {
    dd s = two_sum(a.hi, b.hi);
    double e = s.lo + (a.lo + b.lo);
    return fast_two_sum(s.hi, e);
}

// "Lazy" variants: the pair is kept un-normalised (lo just collects the exact
// rounding errors), which shortens the dependent chain from 10 to 6 fp64 ops.
// dd_renorm() restores |lo| <= ulp(hi)/2 once, at the end of a run of them.
BQ_HD dd dd_acc_lazy(dd a, double v) {
    dd s = two_sum(a.hi, v);
    return dd{s.hi, a.lo + s.lo};
}
BQ_HD dd dd_add_lazy(dd a, dd b) {
    dd s = two_sum(a.hi, b.hi);
    return dd{s.hi, (a.lo + b.lo) + s.lo};
}
BQ_HD dd dd_renorm(dd a) { return two_sum(a.hi, a.lo); }

BQ_HD dd dd_add_d(dd a, double b) {
    dd s = two_sum(a.hi, b);
    double e = s.lo + a.lo;
    return fast_two_sum(s.hi, e);
}

// (a.hi + a.lo) * b rounded to a dd.
BQ_HD dd dd_mul_d(dd a, double b) {
    double p = a.hi * b;
    double e = fma(a.hi, b, -p);
    e = fma(a.lo, b, e);
    return fast_two_sum(p, e);
}

BQ_HD double dd_to_double(dd a) { return a.hi + a.lo; }

// fp32 Kahan step used by the compensated-fp32 instantiation.
BQ_HD void kahan_add(float& sum, float& comp, float v) {
    float y = v - comp;
    float t = sum + y;
    comp = (t - sum) - y;
    sum = t;
}

}  // namespace bq
