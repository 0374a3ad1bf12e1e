// riemann — definite integral of sin(x) on [0, pi] by a left Riemann sum.
//
// Zero-argument behaviour follows the reference program (riemann.cpp:47-101):
// 1e9 steps (riemann.cpp:10), timer started before any set-up (:51), two lines
// on stdout in the reference's formats (:92, :94-96).  The reference needs
// `mpirun -n P` with rank 0 idle (:65-86); here the workers are GPUs whose fused
// kernels combine their partial sums over NVLink, or host threads with
// --backend cpu.
//
// Extra flags (a superset; the reference takes none): --n N|2^k --a A --b B
// --rule left|right|mid|trap --integrand sin|acc|vel|dis|profile --dtype fp64|fp32
// --eval auto|direct|rotate --world 1|2|4|8 --backend auto|cpu --threads T --json
#include <unistd.h>

#include <cmath>
#include <iostream>

#include "cli_common.h"

int main(int argc, char** argv) {
    const double fstart = cli::now_s();
    cli::Args args{argc, argv};
    bq_job job{};
    job.integrand = cli::parse_integrand(args.get("--integrand", "sin"));
    job.rule = cli::parse_rule(args.get("--rule", "left"));
    job.dtype = cli::parse_dtype(args.get("--dtype", "fp64"));
    job.eval = cli::parse_eval(args.get("--eval", "auto"));
    job.a = args.getd("--a", 0.0);
    job.b = args.getd("--b", job.integrand == BQ_F_SIN ? M_PI : 1800.0);
    job.n = args.getu("--n", 1000000000ull);
    const int world = cli::pick_world(args);

    bq_result res{};
    float steady_ms = 0.f;
    // --json also wants the device time of a second, steady-state launch (the first one loads the
    // kernel image); that adds well under a millisecond to the legacy wall-clock line.
    if (cli::run_integrate(job, world, cli::cpu_threads(args), &res, args.has("--json") ? &steady_ms : nullptr) != BQ_OK)
        cli::die("integrate");
    const double fstop = cli::now_s();

    printf("%lf seconds\n", fstop - fstart);
    std::cout.precision(15);
    std::cout << "The integral of f(x) from 0.0 to " << job.b << " with " << double(job.n)
              << " steps is " << res.value << std::endl;
    if (args.has("--json")) {
        const double exact = job.integrand == BQ_F_SIN
                                 ? bq_oracle_sin_sum(job.a, job.b, job.n, job.rule) : NAN;
        printf("{\"program\": \"riemann\", \"n\": %llu, \"world\": %d, \"backend\": \"%s\", "
               "\"value\": %.17g, \"rel_err_vs_oracle\": %.3g, \"device_ms_first_call\": %.6f, "
               "\"device_ms\": %.6f, \"wall_s\": %.6f, \"g_intervals_per_s_device\": %.3f}\n",
               (unsigned long long)job.n, world, world ? "cuda" : "cpu", res.value,
               std::isnan(exact) ? NAN : std::fabs(res.value - exact) / std::fabs(exact),
               res.device_ms, steady_ms > 0 ? steady_ms : res.device_ms, fstop - fstart,
               (steady_ms > 0 ? steady_ms : res.device_ms) > 0
                   ? double(job.n) / ((steady_ms > 0 ? steady_ms : res.device_ms) * 1e-3) / 1e9 : 0.0);
    }
    return 0;
}
