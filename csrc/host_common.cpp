// Host-only pieces: plan construction, oracles, the default table, error text.
#include "host_common.h"

#include "integrands.cuh"

#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

namespace bq {

static const double kDefaultProfile[] = {
#include "profile_table.inc"
};
static_assert(sizeof(kDefaultProfile) / sizeof(double) == 1801, "ex4vel.h has 1801 samples");

const double* default_profile(int* len) {
    if (len) *len = int(sizeof(kDefaultProfile) / sizeof(double));
    return kDefaultProfile;
}

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
const char* last_error() { return g_err; }

// Custom integrands: a table of registered (launch, eval) pairs.  A job names its integrand by
// code (BQ_F_CUSTOM + slot), so every job carries its own kernel: two threads, or two expressions
// used alternately, never see each other's entry points.  Slot 0 is the legacy "active" integrand
// of bq_set_custom_integrand; slots 1.. are write-once.
struct CustomSlot {
    bq_custom_launch_fn launch = nullptr;
    bq_custom_eval_fn eval = nullptr;
};
static std::mutex g_custom_mu;
static CustomSlot g_custom[BQ_MAX_CUSTOM + 1];
static int g_custom_used = 1;

static CustomSlot custom_slot(int integrand) {
    const int slot = integrand - BQ_F_CUSTOM;
    if (slot < 0 || slot > BQ_MAX_CUSTOM) return CustomSlot{};
    std::lock_guard<std::mutex> lock(g_custom_mu);
    return slot < g_custom_used ? g_custom[slot] : CustomSlot{};
}
bq_custom_launch_fn custom_launch(int integrand) { return custom_slot(integrand).launch; }
bq_custom_eval_fn custom_eval(int integrand) { return custom_slot(integrand).eval; }
void set_custom(bq_custom_launch_fn l, bq_custom_eval_fn e) {
    std::lock_guard<std::mutex> lock(g_custom_mu);
    g_custom[0].launch = l, g_custom[0].eval = e;
}
int register_custom(bq_custom_launch_fn l, bq_custom_eval_fn e) {
    std::lock_guard<std::mutex> lock(g_custom_mu);
    for (int i = 1; i < g_custom_used; ++i)
        if (g_custom[i].launch == l && g_custom[i].eval == e) return BQ_F_CUSTOM + i;   // already known
    if (g_custom_used > BQ_MAX_CUSTOM) return -1;
    g_custom[g_custom_used].launch = l, g_custom[g_custom_used].eval = e;
    return BQ_F_CUSTOM + g_custom_used++;
}

// n below which AUTO stays with DIRECT evaluation: the ROTATE set-up (three
// sincos per thread plus one per 32 tiles) has to be paid back.  Measured on B200
// (profiles/v6/sweep_n_v6.jsonl, device us DIRECT / ROTATE): 2^18 14.9 / 15.3,
// 2^20 19.1 / 17.2, 2^22 27.8 / 21.4, 2^24 62 / 25.  A function of n alone, so
// results do not depend on the world size.
static constexpr uint64_t kRotateMinN = 1ull << 20;

int make_plan(const bq_job& job, int rank, int world, uint32_t rows_override, int table_len,
              Plan* P) {
    memset(P, 0, sizeof(*P));
    if (job.n < 1 || job.n > kMaxN) {
        set_error("n=%llu out of range [1, 2^40]", (unsigned long long)job.n);
        return BQ_ERR_INVALID;
    }
    if (world < 1 || world > BQ_MAX_WORLD || (kSegments % world) != 0 || rank < 0 || rank >= world) {
        set_error("world=%d rank=%d: world must be 1, 2, 4 or 8", world, rank);
        return BQ_ERR_INVALID;
    }
    if (job.integrand < 0 || job.integrand > BQ_F_CUSTOM + BQ_MAX_CUSTOM || job.rule < 0 || job.rule > BQ_RULE_TRAP ||
        job.dtype < 0 || job.dtype > BQ_F32C || job.eval < 0 || job.eval > BQ_EVAL_ROTATE) {
        set_error("bad enum in job (integrand=%d rule=%d dtype=%d eval=%d)", job.integrand, job.rule,
                  job.dtype, job.eval);
        return BQ_ERR_INVALID;
    }
    if (!isfinite(job.a) || !isfinite(job.b)) {
        set_error("non-finite interval");
        return BQ_ERR_INVALID;
    }
    P->n = job.n;
    P->a = job.a;
    P->b = job.b;
    P->h = (job.b - job.a) / double(job.n);   // as riemann.cpp:31
    switch (job.rule) {
        case BQ_RULE_RIGHT: P->off = 1.0; break;
        case BQ_RULE_MID: P->off = 0.5; break;
        default: P->off = 0.0; break;          // LEFT; TRAP = LEFT + end correction
    }
    P->integrand = job.integrand;
    P->rule = job.rule;
    P->dtype = job.dtype;
    for (int i = 0; i < 4; ++i) P->p[i] = job.p[i];
    const bool trig = job.integrand == BQ_F_SIN || job.integrand == BQ_F_ACC || job.integrand == BQ_F_VEL ||
                      job.integrand == BQ_F_DIS;
    if (job.integrand >= BQ_F_CUSTOM && !custom_eval(job.integrand) && !custom_launch(job.integrand)) {
        set_error("integrand code %d: no custom integrand is registered under it "
                  "(bq_register_custom_integrand / bq_set_custom_integrand)", job.integrand);
        return BQ_ERR_INVALID;
    }
    if (job.integrand == BQ_F_ACC || job.integrand == BQ_F_VEL || job.integrand == BQ_F_DIS) {
        if (P->p[0] == 0.0) P->p[0] = kTscale;
        if (P->p[1] == 0.0) P->p[1] = kAscale;
        if (P->p[2] == 0.0) P->p[2] = kVscale;
    }
    if (job.integrand == BQ_F_PROFILE) {
        if (P->p[0] == 0.0) P->p[0] = 1.0;
        if (table_len < 2 || table_len > kMaxTable) {
            set_error("PROFILE table length %d out of range [2, %d]", table_len, kMaxTable);
            return BQ_ERR_INVALID;
        }
        P->table_len = uint32_t(table_len);
        const double xmax = double(table_len - 1) * P->p[0];
        const double lo = job.a < job.b ? job.a : job.b, hi = job.a < job.b ? job.b : job.a;
        if (lo < 0.0 || hi > xmax * (1.0 + 1e-15)) {
            set_error("PROFILE integrand is defined on [0, %g], got [%g, %g]", xmax, lo, hi);
            return BQ_ERR_RANGE;
        }
    }
    P->omega = (trig && job.integrand != BQ_F_SIN) ? 1.0 / P->p[0] : 1.0;
    if (job.eval == BQ_EVAL_AUTO)
        P->eval = (trig && job.n >= kRotateMinN) ? BQ_EVAL_ROTATE : BQ_EVAL_DIRECT;
    else
        P->eval = (job.eval == BQ_EVAL_ROTATE && trig) ? BQ_EVAL_ROTATE : BQ_EVAL_DIRECT;

    const uint64_t total_rows = (job.n + kWidth - 1) / kWidth;
    uint64_t rows = (total_rows + kParts - 1) / kParts;
    if (rows_override) rows = rows_override;
    if (rows < 1) rows = 1;
    if (rows > kMaxRows) rows = kMaxRows;
    P->rows = uint32_t(rows);
    P->tile_pts = rows * kWidth;
    P->n_tiles = (job.n + P->tile_pts - 1) / P->tile_pts;

    P->rank = rank;
    P->world = world;
    const uint32_t segs = kSegments / world;
    P->seg_lo = rank * segs;
    P->seg_hi = P->seg_lo + segs;
    P->chunk_lo = P->seg_lo * kChunksPerSegment;
    P->chunk_hi = P->seg_hi * kChunksPerSegment;
    P->flags = 0;
    // DIRECT evaluation of a trig integrand may use the bounded-argument sin/cos when the whole
    // interval (plus one step of slack for the rule offset) keeps |angle| inside its range.
    if (trig) {
        const double w = fabs(P->omega);
        const double amax = (fabs(job.a) > fabs(job.b) ? fabs(job.a) : fabs(job.b)) + fabs(P->h);
        if (w * amax <= kFastTrigMax) {
            P->flags |= kFlagFastTrig;
            if (w * fabs(P->h) * double(kWidth) <= 1.0 / 1024.0) P->flags |= kFlagRowK;
        }
        if (w * fabs(P->h) * double(kWidth) <= 1.0 / 256.0) P->flags |= kFlagFineGrid;
    }
    P->timeout_ms = 2000;
    return BQ_OK;
}

void fill_row_consts(const Plan& P, RowConsts* out) {
    // angle(x) = x for sin, x / tscale for the train models (integrands.cuh)
    const bool scaled = P.integrand != BQ_F_SIN;
    const double tscale = scaled ? P.p[0] : 1.0;
    struct Cache {
        bool valid = false;
        double h = 0.0, tscale = 0.0;
        RowConsts rc;
    };
    static thread_local Cache cache;
    if (!(cache.valid && cache.h == P.h && cache.tscale == tscale)) {
        for (int r = 0; r < kFlushTiles * kMaxRows; ++r) {
            const double x = double(r) * double(kWidth) * P.h;
            double sn, cs;
            ::sincos(scaled ? x / tscale : x, &sn, &cs);
            cache.rc.cs[r][0] = cs;
            cache.rc.cs[r][1] = sn;
        }
        cache.h = P.h, cache.tscale = tscale, cache.valid = true;
    }
    *out = cache.rc;
}

// Closed form of sum_{i<n} sin(a + (i+off) h) = sin(a + off h + (n-1)h/2) sin(nh/2) / sin(h/2),
// evaluated in long double (x87 80-bit on the hosts this runs on).
long double oracle_sin_sum(double a, double b, uint64_t n, int rule) {
    const long double A = a, B = b, N = (long double)n;
    const long double h = (B - A) / N;
    long double off = 0.0L;
    if (rule == BQ_RULE_RIGHT) off = 1.0L;
    if (rule == BQ_RULE_MID) off = 0.5L;
    long double sum;
    if (h == 0.0L) {
        sum = N * sinl(A);
    } else {
        sum = sinl(A + off * h + (N - 1.0L) * h * 0.5L) * sinl(N * h * 0.5L) / sinl(h * 0.5L);
    }
    long double val = h * sum;
    if (rule == BQ_RULE_TRAP) val += 0.5L * h * (sinl(B) - sinl(A));
    return val;
}

}  // namespace bq

extern "C" {

const char* bq_version(void) { return "b200quad 0.1 (sm_100a)"; }
int bq_set_custom_integrand(bq_custom_launch_fn launch, bq_custom_eval_fn eval) {
    bq::set_custom(launch, eval);
    return BQ_OK;
}
int bq_register_custom_integrand(bq_custom_launch_fn launch, bq_custom_eval_fn eval) {
    if (!launch && !eval) return -1;
    const int code = bq::register_custom(launch, eval);
    if (code < 0) bq::set_error("all %d custom integrand slots are in use", BQ_MAX_CUSTOM);
    return code;
}
const char* bq_last_error(void) { return bq::last_error(); }
const double* bq_default_profile(int* len) { return bq::default_profile(len); }
size_t bq_sizeof_params(void) { return sizeof(bq::Plan) + sizeof(bq::RowConsts); }
double bq_sin_bounded(double x) { return bq::sincos_bounded<false>(x); }
double bq_cos_bounded(double x) { return bq::sincos_bounded<true>(x); }
double bq_oracle_sin_sum(double a, double b, uint64_t n, int rule) {
    return (double)bq::oracle_sin_sum(a, b, n, rule);
}

}  // extern "C"
