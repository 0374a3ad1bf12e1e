// train4 — the "train problem" of 4main.c: sample the velocity profile at 10000
// steps per second, running sum (position), running sum of the running sum.
//
// Zero-argument behaviour follows the reference (4main.c:42-247): the lines
// "Step size of 10000" (:73), "%lf seconds" (:239) and "Total distance
// traveled = %lf" computed from element N-2 of the first table (:241).  The
// reference splits the table over MPI ranks, gathers the slices on rank 0, fixes
// the carries serially and broadcasts 144 MB twice (:141-157,205-221); here one
// fused kernel per GPU does both running sums in one pass and, with several
// GPUs, writes every tile into every GPU's table (--no-replicate keeps slices).
#include <unistd.h>

#include <cmath>
#include <vector>

#include "cli_common.h"

int main(int argc, char** argv) {
    cli::Args args{argc, argv};
    const uint64_t steps_per_sec = args.getu("--steps-per-sec", 10000);
    const uint64_t seconds = args.getu("--seconds", 1800);
    bq_scan_job job{};
    job.integrand = cli::parse_integrand(args.get("--integrand", "profile"));
    job.replicate = args.has("--no-replicate") ? 0 : 1;
    job.x0 = 0.0;
    job.dx = 1.0 / double(steps_per_sec);
    job.n = seconds * steps_per_sec;
    const int world = cli::pick_world(args);
    const int threads = cli::cpu_threads(args);

    bq_scan_result res{};
    float device_ms = 0.f;
    // As in the reference, table storage is set up before the timer starts (4main.c:52-67).
    if (world == 0) {
        std::vector<double> P(job.n), Q(job.n);
        const double fstart = cli::now_s();
        printf("Step size of %ld\n", long(steps_per_sec));
        if (bq_cpu_cumulative(&job, threads, P.data(), Q.data(), nullptr, &res) != BQ_OK)
            cli::die("cumulative");
        printf("%lf seconds\n", cli::now_s() - fstart);
    } else {
        bq_world* w = nullptr;
        if (bq_world_create(world, nullptr, &w) != BQ_OK) cli::die("world");
        double *dP[BQ_MAX_WORLD] = {}, *dQ[BQ_MAX_WORLD] = {};
        const uint64_t tiles = (job.n + 2047) / 2048;
        for (int r = 0; r < world; ++r) {
            uint64_t len = job.n;
            if (!job.replicate && world > 1) {
                const uint64_t lo = (uint64_t(r) * (8 / world) * tiles / 8) * 2048;
                uint64_t hi = (uint64_t(r + 1) * (8 / world) * tiles / 8) * 2048;
                if (hi > job.n) hi = job.n;
                len = hi - lo;
            }
            const int dev = bq_ctx_device(bq_world_ctx(w, r));
            if (bq_device_alloc(dev, len * 8, (void**)&dP[r]) != BQ_OK ||
                bq_device_alloc(dev, len * 8, (void**)&dQ[r]) != BQ_OK)
                cli::die("device_alloc");
        }
        const double fstart = cli::now_s();
        printf("Step size of %ld\n", long(steps_per_sec));
        bq_scan_result all[BQ_MAX_WORLD];
        if (bq_world_cumulative(w, &job, dP, dQ, all) != BQ_OK) cli::die("cumulative");
        printf("%lf seconds\n", cli::now_s() - fstart);
        res = all[0];
        for (int r = 0; r < world; ++r) device_ms = all[r].device_ms > device_ms ? all[r].device_ms : device_ms;
        for (int r = 0; r < world; ++r) {
            const int dev = bq_ctx_device(bq_world_ctx(w, r));
            bq_device_free(dev, dP[r]);
            bq_device_free(dev, dQ[r]);
        }
        bq_world_destroy(w);
    }
    printf("Total distance traveled = %lf\n", res.p_second_last / double(steps_per_sec));
    if (args.has("--json"))
        printf("{\"program\": \"train4\", \"n\": %llu, \"world\": %d, \"p_last\": %.17g, "
               "\"p_second_last\": %.17g, \"q_last\": %.17g, \"device_ms\": %.6f}\n",
               (unsigned long long)job.n, world, res.p_last, res.p_second_last, res.q_last, device_ms);
    return 0;
}
