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
    template <class R>
    BQ_HD R eval(R x) const { return bq_sin(x); }
    BQ_HD double angle(double x) const { return x; }
    template <class R>
    BQ_HD R eval_sc(R, R s, R) const { return s; }
};

struct AccF {
    static constexpr bool kFusedSum = true;     // -ascale * sin
    static constexpr bool kTrig = true, kNeedS = true, kNeedC = false, kNeedX = false;
    double tscale, ascale;
    BQ_HD double fuse_alpha() const { return -ascale; }
    BQ_HD double fuse_beta() const { return 0.0; }
    BQ_HD double fuse_gamma() const { return 0.0; }
    template <class R>
    BQ_HD R eval(R x) const { return -(bq_sin(x / R(tscale)) * R(ascale)); }
    BQ_HD double angle(double x) const { return x / tscale; }
    template <class R>
    BQ_HD R eval_sc(R, R s, R) const { return -(s * R(ascale)); }
};

struct VelF {
    static constexpr bool kFusedSum = true;     // vscale - vscale * cos: the constant is added once, as gamma * n
    static constexpr bool kTrig = true, kNeedS = false, kNeedC = true, kNeedX = false;
    double tscale, vscale;
    BQ_HD double fuse_alpha() const { return 0.0; }
    BQ_HD double fuse_beta() const { return -vscale; }
    BQ_HD double fuse_gamma() const { return vscale; }
    template <class R>
    BQ_HD R eval(R x) const { return (-bq_cos(x / R(tscale)) + R(1)) * R(vscale); }
    BQ_HD double angle(double x) const { return x / tscale; }
    template <class R>
    BQ_HD R eval_sc(R, R, R c) const { return (-c + R(1)) * R(vscale); }
};

struct DisF {
    static constexpr bool kFusedSum = false;
    BQ_HD double fuse_alpha() const { return 0.0; }
    BQ_HD double fuse_beta() const { return 0.0; }
    BQ_HD double fuse_gamma() const { return 0.0; }
    static constexpr bool kTrig = true, kNeedS = true, kNeedC = false, kNeedX = true;
    double tscale, vscale;
    template <class R>
    BQ_HD R eval(R x) const { return R(vscale) * bq_fma(-R(tscale), bq_sin(x / R(tscale)), x); }
    BQ_HD double angle(double x) const { return x / tscale; }
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
    BQ_HD double angle(double) const { return 0.0; }
    template <class R>
    BQ_HD R eval_sc(R x, R, R) const { return eval(x); }
};

}  // namespace bq
