// Host-only helpers shared by the CUDA runtime, the CPU backend and the CLIs.
#pragma once
#include <stdint.h>

#include "plan.h"

namespace bq {

const double* default_profile(int* len);
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
const char* last_error();
long double oracle_sin_sum(double a, double b, uint64_t n, int rule);

// The active custom integrand (bq_set_custom_integrand): null until one is registered.
bq_custom_launch_fn custom_launch();
bq_custom_eval_fn custom_eval();
void set_custom(bq_custom_launch_fn launch, bq_custom_eval_fn eval);

}  // namespace bq
