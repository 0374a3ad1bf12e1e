// Shared by the three reference-compatible command-line programs.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <string>

#include "b200quad.h"

namespace cli {

inline double now_s() {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);   // the reference's clock (riemann.cpp:51, 4main.c:67)
    return double(t.tv_sec) + double(t.tv_nsec) / 1e9;
}

struct Args {
    int argc;
    char** argv;
    bool has(const char* flag) const {
        for (int i = 1; i < argc; ++i)
            if (!strcmp(argv[i], flag)) return true;
        return false;
    }
    const char* get(const char* flag, const char* dflt) const {
        for (int i = 1; i + 1 < argc; ++i)
            if (!strcmp(argv[i], flag)) return argv[i + 1];
        return dflt;
    }
    double getd(const char* flag, double dflt) const {
        const char* v = get(flag, nullptr);
        return v ? strtod(v, nullptr) : dflt;
    }
    uint64_t getu(const char* flag, uint64_t dflt) const {
        const char* v = get(flag, nullptr);
        if (!v) return dflt;
        if (!strncmp(v, "2^", 2)) return 1ull << atoi(v + 2);
        return uint64_t(strtod(v, nullptr));
    }
};

inline int parse_enum(const char* v, const char* const* names, int n, const char* what) {
    for (int i = 0; i < n; ++i)
        if (!strcmp(v, names[i])) return i;
    fprintf(stderr, "unknown %s '%s'\n", what, v);
    exit(2);
}

inline int parse_integrand(const char* v) {
    static const char* const k[] = {"sin", "acc", "vel", "dis", "profile"};
    return parse_enum(v, k, 5, "integrand");
}
inline int parse_rule(const char* v) {
    static const char* const k[] = {"left", "right", "mid", "trap"};
    return parse_enum(v, k, 4, "rule");
}
inline int parse_dtype(const char* v) {
    static const char* const k[] = {"fp64", "fp32"};
    return parse_enum(v, k, 2, "dtype");
}
inline int parse_eval(const char* v) {
    static const char* const k[] = {"auto", "direct", "rotate"};
    return parse_enum(v, k, 3, "eval");
}

// Default world of the zero-argument programs: ONE GPU when there is one (0 = CPU backend).
// The programs print the reference's wall-clock line, which spans context creation to the result
// (riemann.cpp:51,91-92; cintegrate.cu:104,139-140): the reference's problem sizes (1e9 intervals,
// 18e6 samples) are 0.15 ms / 0.06 ms of work on one B200, while every extra GPU adds a context
// creation and peer wiring (tens of milliseconds each) to that line — the fastest time-to-solution
// is a world of 1 (measured: profiles/r2/cli_wall.jsonl).  --world 2|4|8 spreads larger problems.
inline int default_world() { return bq_cuda_device_count() >= 1 ? 1 : 0; }

inline int pick_world(const Args& a) {
    if (!strcmp(a.get("--backend", "auto"), "cpu")) return 0;
    const int dflt = default_world();
    const int w = int(a.getu("--world", uint64_t(dflt)));
    if (w > 0 && bq_cuda_device_count() < w) {
        fprintf(stderr, "--world %d but only %d CUDA device(s)\n", w, bq_cuda_device_count());
        exit(2);
    }
    return w;
}

inline int cpu_threads(const Args& a) {
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return int(a.getu("--threads", uint64_t(n > 0 ? n : 1)));
}

// integrate on `world` GPUs (0 = CPU backend); returns rank 0's result.  The first launch of a
// process also pays for loading the kernel image; with `steady_ms` the same job is launched once
// more and *steady_ms gets that second launch's device time (max over ranks) — what --json reports
// next to the first-call wall time of the legacy line.
inline int run_integrate(const bq_job& job, int world, int threads, bq_result* out, float* steady_ms = nullptr) {
    if (world == 0) return bq_cpu_integrate(&job, threads, out);
    bq_world* w = nullptr;
    int rc = bq_world_create(world, nullptr, &w);
    if (rc != BQ_OK) return rc;
    bq_result res[BQ_MAX_WORLD];
    rc = bq_world_integrate(w, &job, res);
    *out = res[0];
    float ms = 0.f;
    for (int r = 0; r < world; ++r) ms = res[r].device_ms > ms ? res[r].device_ms : ms;
    out->device_ms = ms;
    if (rc == BQ_OK && steady_ms) {
        bq_result again[BQ_MAX_WORLD];
        if (bq_world_integrate(w, &job, again) == BQ_OK) {
            float m2 = 0.f;
            for (int r = 0; r < world; ++r) m2 = again[r].device_ms > m2 ? again[r].device_ms : m2;
            *steady_ms = m2;
        }
    }
    bq_world_destroy(w);
    return rc;
}

inline void die(const char* what) {
    fprintf(stderr, "%s: %s\n", what, bq_last_error());
    exit(1);
}

}  // namespace cli
