// Integrand library (host + device functors).
//
//  SinF      sin(x)                         riemann.cpp:37, cintegrate.cu:68
//  AccF      -sin(x/tscale)*ascale          riemann.cpp:103-106
//  VelF      (1-cos(x/tscale))*vscale       riemann.cpp:108-111
//  DisF      vscale*(x-tscale*sin(x/tscale)) riemann.cpp:113-116
//  ProfileF  lerp of a uniform table        4main.c:249-269, cintegrate.cu:23-44
//
// Each functor offers
//   eval(x)            - plain evaluation (DIRECT mode; any x)
//   angle(x)           - for trig functors, the argument of sin/cos
//   eval_sc(x, s, c)   - evaluation given s=sin(angle(x)), c=cos(angle(x))
//                        (ROTATE mode: the kernel supplies s/c by rotation)
// templated on the real type so the compensated-fp32 path shares them.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "bq_dd.cuh"

namespace bq {

BQ_HD double bq_fma(double a, double b, double c) { return fma(a, b, c); }
BQ_HD float bq_fma(float a, float b, float c) { return fmaf(a, b, c); }
BQ_HD double bq_sin(double x) { return sin(x); }
BQ_HD double bq_cos(double x) { return cos(x); }
BQ_HD float bq_sin(float x) { return sinf(x); }
BQ_HD float bq_cos(float x) { return cosf(x); }

// ---- bounded-argument sine / cosine -------------------------------------------------------
// libm / libdevice sin() carries a Payne-Hanek path for huge arguments and, on the GPU, costs
// ~40 fp64-pipe instructions per call.  A definite integral knows its argument range before the
// first sample is taken (the plan checks |angle| <= kFastTrigMax over [a, b]), so DIRECT
// evaluation can use this instead: k = rint(x * 2/pi) through the conversion unit (not the fp64
// pipe), a two-constant Cody-Waite reduction (k * PIO2_HI is exact for |k| < 2^20: PIO2_HI has 33
// significant bits), and ONE degree-6 polynomial in r^2 whose coefficients are selected by the
// quadrant (fdlibm's kernel_sin / kernel_cos minimax sets, error < 2^-57 on |r| <= pi/4):
//     sin(r) = r + (r t)(S1 + S2 t + .. + S6 t^5)      cos(r) = 1 + t(-1/2 + C1 t + .. + C6 t^6)
// 12 fp64 operations, no branch, identical bits on host and device (same code, explicit fma).
// Measured against long-double sinl/cosl over |x| <= 1e6: absolute error <= 1.11 * 2^-53, relative
// < 1.4 ulp away from the zeros (tests/test_cpu_quad.py); next to a zero the error is absolute
// (|x| * 2^-87), which is what a sum of samples needs.
constexpr double kFastTrigMax = 1.0e6;

namespace trig {
constexpr double kTwoOverPi = 6.36619772367581382433e-01;
constexpr double kPio2Hi = 1.57079632673412561417e+00;     // first 33 bits of pi/2
constexpr double kPio2Lo = 6.07710050650619224932e-11;     // pi/2 - kPio2Hi
constexpr double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
                 S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
                 S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
constexpr double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
                 C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
                 C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
}  // namespace trig

// Reduced argument r in [-pi/4, pi/4] and quadrant q (already advanced by one for a cosine):
// the wanted value is (-1)^(q>>1) * (q odd ? cos r : sin r).
struct TrigRed {
    double r;
    int q;
    double kd;      // the multiple of pi/2 that was taken off, as a double
};
// Reduction with a given multiple (the samples of one row of one thread are a tiny angle apart and
// share the first one's: |r| may then exceed pi/4 by the row's span, which the plan keeps below
// 2^-10 rad — far inside the range where the polynomials hold their accuracy).
template <bool COS>
BQ_HD TrigRed trig_reduce_with(double x, const TrigRed& first) {
    double r = fma(-first.kd, trig::kPio2Hi, x);
    r = fma(-first.kd, trig::kPio2Lo, r);
    return TrigRed{r, first.q, first.kd};
}
// BQ_TRIG_MAGIC=1: k through the 1.5 * 2^52 trick (one fma + one add on the fp64 pipe, the integer
// falls out of the low word) instead of a conversion pair (F2I.F64 + I2F.F64 on the conversion unit).
// Measured on B200 at n = 2^32 (DIRECT sin, fp64): conversions 883 G/s, magic 955 G/s.
#ifndef BQ_TRIG_MAGIC
#define BQ_TRIG_MAGIC 1
#endif
template <bool COS>
BQ_HD TrigRed trig_reduce(double x) {
#if BQ_TRIG_MAGIC
    constexpr double kMagic = 6755399441055744.0;               // 1.5 * 2^52: ulp = 1, so the fma rounds to an integer
    const double tm = fma(x, trig::kTwoOverPi, kMagic);
    const double kd = tm - kMagic;
    long long bits;
    memcpy(&bits, &tm, sizeof(bits));
    const int k = int(uint32_t(uint64_t(bits)));                // low 32 bits of the mantissa = k (two's complement)
#else
#if defined(__CUDA_ARCH__)
    const int k = __double2int_rn(x * trig::kTwoOverPi);
#else
    const int k = int(nearbyint(x * trig::kTwoOverPi));         // round-to-nearest-even, as cvt.rni
#endif
    const double kd = double(k);
#endif
    double r = fma(-kd, trig::kPio2Hi, x);                      // exact
    r = fma(-kd, trig::kPio2Lo, r);
    return TrigRed{r, k + (COS ? 1 : 0), kd};
}
// The three forms below give the same bits for the same (r, parity): the selected form starts the
// sine's Horner chain with fma(0, t, S6) = S6.  The kernel takes a branch-free unselected form when
// a whole warp-row has one parity (nearly always: neighbouring samples are a tiny angle apart) and
// the selected form otherwise; the CPU tree always takes the selected form.
BQ_HD double trig_poly_sin(double r) {
    using namespace trig;
    const double t = r * r;
    double p = fma(S6, t, S5);
    p = fma(p, t, S4);
    p = fma(p, t, S3);
    p = fma(p, t, S2);
    p = fma(p, t, S1);
    return fma(r * t, p, r);
}
BQ_HD double trig_poly_cos(double r) {
    using namespace trig;
    const double t = r * r;
    double p = fma(C6, t, C5);
    p = fma(p, t, C4);
    p = fma(p, t, C3);
    p = fma(p, t, C2);
    p = fma(p, t, C1);
    p = fma(p, t, -0.5);
    return fma(t, p, 1.0);
}
BQ_HD double trig_poly_sel(double r, bool odd) {
    using namespace trig;
    const double t = r * r;
    double p = odd ? C6 : 0.0;
    p = fma(p, t, odd ? C5 : S6);
    p = fma(p, t, odd ? C4 : S5);
    p = fma(p, t, odd ? C3 : S4);
    p = fma(p, t, odd ? C2 : S3);
    p = fma(p, t, odd ? C1 : S2);
    p = fma(p, t, odd ? -0.5 : S1);
    const double rt = r * t;
    return fma(odd ? t : rt, p, odd ? 1.0 : r);
}
BQ_HD double trig_sign(double v, int q) { return (q & 2) ? -v : v; }

template <bool COS>
BQ_HD double sincos_bounded(double x) {
    const TrigRed red = trig_reduce<COS>(x);
    return trig_sign(trig_poly_sel(red.r, (red.q & 1) != 0), red.q);
}
// eval_fast of a trig functor in fp64: the bounded sin (or cos) of angle_fast(x) fed to eval_sc —
// exactly what the row evaluator of the kernel computes (RowDirect, quad_core.cuh), sample by sample.
template <class F>
BQ_HD double trig_eval_fast(const F& f, double x) {
    constexpr bool COS = F::kNeedC;
    const double sc = sincos_bounded<COS>(f.angle_fast(x));
    return COS ? f.template eval_sc<double>(x, 0.0, sc) : f.template eval_sc<double>(x, sc, 0.0);
}

struct SinF {
    static constexpr bool kTrig = true, kNeedS = true, kNeedC = false, kNeedX = false;
    // ROTATE/fp64: the sample sin(x0 + dx) = S*cj + C*sj is added to the running sum by two
    // chained FMAs (acc = fma(C, sj, fma(S, cj, acc))) instead of being rounded to a double first
    // and added: 2 fp64 instructions per interval instead of 3, one rounding fewer per term.
    static constexpr bool kFusedSum = true;
    // f = alpha * sin(angle) + beta * cos(angle) + gamma (what the fused loop needs, see fused_consts)
    BQ_HD double fuse_alpha() const { return 1.0; }
    BQ_HD double fuse_beta() const { return 0.0; }
    BQ_HD double fuse_gamma() const { return 0.0; }
    BQ_HD double fuse_delta() const { return 0.0; }
    BQ_HD double fuse_scale() const { return 1.0; }
    template <class R>
    BQ_HD R eval(R x) const { return bq_sin(x); }
    BQ_HD double eval_fast_d(double x) const { return trig_eval_fast(*this, x); }   // |x| <= kFastTrigMax (plan)
    template <class R>
    BQ_HD R eval_fast(R x) const {
        if constexpr (sizeof(R) == 8) return eval_fast_d(x);
        return eval<R>(x);
    }
    BQ_HD double angle(double x) const { return x; }
    BQ_HD double angle_fast(double x) const { return x; }
    template <class R>
    BQ_HD R eval_sc(R, R s, R) const { return s; }
};

struct AccF {
    static constexpr bool kFusedSum = true;     // -ascale * sin
    static constexpr bool kTrig = true, kNeedS = true, kNeedC = false, kNeedX = false;
    double tscale, ascale, inv_tscale;
    BQ_HD double fuse_alpha() const { return -ascale; }
    BQ_HD double fuse_beta() const { return 0.0; }
    BQ_HD double fuse_gamma() const { return 0.0; }
    BQ_HD double fuse_delta() const { return 0.0; }
    BQ_HD double fuse_scale() const { return 1.0; }
    template <class R>
    BQ_HD R eval(R x) const { return -(bq_sin(x / R(tscale)) * R(ascale)); }
    BQ_HD double eval_fast_d(double x) const { return trig_eval_fast(*this, x); }
    template <class R>
    BQ_HD R eval_fast(R x) const {
        if constexpr (sizeof(R) == 8) return eval_fast_d(x);
        return eval<R>(x);
    }
    BQ_HD double angle(double x) const { return x / tscale; }
    BQ_HD double angle_fast(double x) const { return x * inv_tscale; }   // one multiply instead of a division
    template <class R>
    BQ_HD R eval_sc(R, R s, R) const { return -(s * R(ascale)); }
};

struct VelF {
    static constexpr bool kFusedSum = true;     // vscale - vscale * cos: the constant is added once, as gamma * n
    static constexpr bool kTrig = true, kNeedS = false, kNeedC = true, kNeedX = false;
    double tscale, vscale, inv_tscale;
    BQ_HD double fuse_alpha() const { return 0.0; }
    BQ_HD double fuse_beta() const { return -vscale; }
    BQ_HD double fuse_gamma() const { return vscale; }
    BQ_HD double fuse_delta() const { return 0.0; }
    BQ_HD double fuse_scale() const { return 1.0; }
    template <class R>
    BQ_HD R eval(R x) const { return (-bq_cos(x / R(tscale)) + R(1)) * R(vscale); }
    BQ_HD double eval_fast_d(double x) const { return trig_eval_fast(*this, x); }
    template <class R>
    BQ_HD R eval_fast(R x) const {
        if constexpr (sizeof(R) == 8) return eval_fast_d(x);
        return eval<R>(x);
    }
    BQ_HD double angle(double x) const { return x / tscale; }
    BQ_HD double angle_fast(double x) const { return x * inv_tscale; }   // one multiply instead of a division
    template <class R>
    BQ_HD R eval_sc(R, R, R c) const { return (-c + R(1)) * R(vscale); }
};

struct DisF {
    // vscale * x - vscale * tscale * sin(x / tscale): the sine part goes through the fused loop sample by
    // sample; the linear part, like a constant part, is summed in closed form over the grid
    // (delta * sum of the x_i, an arithmetic series in dd arithmetic: finish_result)
    // f = scale * (alpha * sin + delta * x) with alpha = -tscale, delta = 1, scale = vscale: no product of
    // two parameters is rounded before the two sums meet (x and tscale * sin cancel to x^3 / (6 tscale^2)
    // for small x; a rounded vscale * tscale cost 4e-12 there)
    static constexpr bool kFusedSum = true;
    BQ_HD double fuse_alpha() const { return -tscale; }
    BQ_HD double fuse_beta() const { return 0.0; }
    BQ_HD double fuse_gamma() const { return 0.0; }
    BQ_HD double fuse_delta() const { return 1.0; }
    BQ_HD double fuse_scale() const { return vscale; }
    static constexpr bool kTrig = true, kNeedS = true, kNeedC = false, kNeedX = false;
    double tscale, vscale, inv_tscale;
    template <class R>
    BQ_HD R eval(R x) const { return R(vscale) * bq_fma(-R(tscale), bq_sin(x / R(tscale)), x); }
    BQ_HD double eval_fast_d(double x) const { return trig_eval_fast(*this, x); }
    template <class R>
    BQ_HD R eval_fast(R x) const {
        if constexpr (sizeof(R) == 8) return eval_fast_d(x);
        return eval<R>(x);
    }
    BQ_HD double angle(double x) const { return x / tscale; }
    BQ_HD double angle_fast(double x) const { return x * inv_tscale; }   // one multiply instead of a division
    template <class R>
    BQ_HD R eval_sc(R x, R s, R) const { return R(vscale) * bq_fma(-R(tscale), s, x); }
};

// Uniform-grid table, linear interpolation.  The reference reads
// table[(int)t + 1] unconditionally (4main.c:264-268), which runs one past the
// end at t == 1800; here the cell index is clamped to [0, len-2] so the closed
// right end is valid, and anything outside [0, (len-1)*spacing] extrapolates
// the edge cell linearly (the kernel flags it as BQ_ERR_RANGE).
struct ProfileF {
    static constexpr bool kFusedSum = false;
    BQ_HD double fuse_alpha() const { return 0.0; }
    BQ_HD double fuse_beta() const { return 0.0; }
    BQ_HD double fuse_gamma() const { return 0.0; }
    static constexpr bool kTrig = false, kNeedS = false, kNeedC = false, kNeedX = true;
    const double* table;
    int len;
    double inv_dx;
    double step_du;      // table-index distance between consecutive samples of a run (see eval_run)
    BQ_HD int cell_of(double u) const {
        int i = (int)u;
        return i < 0 ? 0 : (i > len - 2 ? len - 2 : i);
    }
    // N samples at x_first + j * step (j = 0..N-1, step = step_du table cells): when the first
    // and the last one lie in the same cell — a run is short against the table spacing, so
    // nearly always — the cell index and the slope are worked out once and every sample is ONE
    // fma, v_j = fma(slope * step_du, j, v_0): the lerp is affine inside a cell.  Otherwise each
    // sample goes through eval().  x_last is the caller's own x of sample N-1 (same rounding as
    // eval would see).  Replaces the per-sample table_accel x3 + faccel of 4main.c:249-269.
    template <int N>
    BQ_HD bool eval_run(double x_first, double x_last, double (&v)[N]) const {
        const double u0 = x_first * inv_dx, u1 = x_last * inv_dx;
        const int i0 = cell_of(u0), i1 = cell_of(u1);
        if (i0 != i1) return false;
        const double t0 = table[i0];
        const double slope = table[i0 + 1] - t0;
        const double v0 = fma(slope, u0 - double(i0), t0);
        const double sh = slope * step_du;
#pragma unroll
        for (int j = 0; j < N; ++j) v[j] = fma(sh, double(j), v0);
        return true;
    }
    // The same run in fp32 (dtype fp32): the cell test and the position inside the cell in fp64
    // (an fp32 x cannot tell neighbouring samples of a fine grid apart), the N lerps as one fp32 fma each.
    template <int N>
    BQ_HD bool eval_run(double x_first, double x_last, float (&v)[N]) const {
        const double u0 = x_first * inv_dx, u1 = x_last * inv_dx;
        const int i0 = cell_of(u0), i1 = cell_of(u1);
        if (i0 != i1) return false;
        const float t0 = float(table[i0]);
        const float slope = float(table[i0 + 1]) - t0;
        const float v0 = bq_fma(slope, float(u0 - double(i0)), t0);
        const float sh = slope * float(step_du);
#pragma unroll
        for (int j = 0; j < N; ++j) v[j] = bq_fma(sh, float(j), v0);
        return true;
    }
    template <class R>
    BQ_HD R eval_fast(R x) const { return eval(x); }
    template <class R>
    BQ_HD R eval(R x) const {
        // (int)u / R(i): measured on B200, these two conversions are cheaper than any
        // conversion-free floor built from fp64 adds (the fp64 pipe is the bound here).
        R u = x * R(inv_dx);
        int i = (int)u;
        i = i < 0 ? 0 : (i > len - 2 ? len - 2 : i);
        R frac = u - R(i);
        R t0 = R(table[i]);
        R t1 = R(table[i + 1]);
        return bq_fma(t1 - t0, frac, t0);
    }
    BQ_HD double angle(double) const { return 0.0; }
    template <class R>
    BQ_HD R eval_sc(R x, R, R) const { return eval(x); }
};

// Host-side stand-in for a custom integrand (BQ_F_CUSTOM): calls the eval entry point of the
// user's translation unit.  The GPU side never sees this type — there the expression itself is
// compiled into the kernel (b200quad/models/jit.py generates that unit).
struct CustomHostF {
    static constexpr bool kFusedSum = false;
    BQ_HD double fuse_alpha() const { return 0.0; }
    BQ_HD double fuse_beta() const { return 0.0; }
    BQ_HD double fuse_gamma() const { return 0.0; }
    static constexpr bool kTrig = false, kNeedS = false, kNeedC = false, kNeedX = true;
    double (*fn)(double, const double*);
    double p[4];
    template <class R>
    BQ_HD R eval(R x) const { return R(fn(double(x), p)); }
    template <class R>
    BQ_HD R eval_fast(R x) const { return eval(x); }
    BQ_HD double angle(double) const { return 0.0; }
    template <class R>
    BQ_HD R eval_sc(R x, R, R) const { return eval(x); }
};

}  // namespace bq
