// Host-only helpers shared by the CUDA runtime, the CPU backend and the CLIs.
#pragma once
#include <stdint.h>

#include "plan.h"

namespace bq {

const double* default_profile(int* len);
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
const char* last_error();
long double oracle_sin_sum(double a, double b, uint64_t n, int rule);

// Custom integrands by integrand code (BQ_F_CUSTOM + slot): null when nothing is registered there.
bq_custom_launch_fn custom_launch(int integrand);
bq_custom_eval_fn custom_eval(int integrand);
void set_custom(bq_custom_launch_fn launch, bq_custom_eval_fn eval);      // slot 0
int register_custom(bq_custom_launch_fn launch, bq_custom_eval_fn eval);  // -> integrand code, -1 when full

}  // namespace bq
