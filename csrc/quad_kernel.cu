// Instantiations of the fused definite-integral kernel (quad_kernel.cuh) for the integrand
// library, and the dispatch on (integrand, dtype, evaluation mode).
#include "quad_kernel.cuh"

#include "host_common.h"

namespace bq {

template <class F>
static cudaError_t launch_f(const Plan& P, const DeviceState& st, const LaunchConfig& cfg,
                            cudaStream_t stream) {
    const bool rot = P.eval == BQ_EVAL_ROTATE;
    if constexpr (F::kTrig) {
        if (P.dtype == BQ_F64)
            return rot ? launch_one<F, double, BQ_EVAL_ROTATE>(P, st, cfg, stream)
                       : launch_one<F, double, BQ_EVAL_DIRECT>(P, st, cfg, stream);
        return rot ? launch_one<F, float, BQ_EVAL_ROTATE>(P, st, cfg, stream)
                   : launch_one<F, float, BQ_EVAL_DIRECT>(P, st, cfg, stream);
    } else {
        if (P.dtype == BQ_F64) return launch_one<F, double, BQ_EVAL_DIRECT>(P, st, cfg, stream);
        return launch_one<F, float, BQ_EVAL_DIRECT>(P, st, cfg, stream);
    }
}

cudaError_t launch_quad(const Plan& P, const DeviceState& st, const LaunchConfig& cfg,
                        cudaStream_t stream) {
    switch (P.integrand) {
        case BQ_F_SIN: return launch_f<SinF>(P, st, cfg, stream);
        case BQ_F_ACC: return launch_f<AccF>(P, st, cfg, stream);
        case BQ_F_VEL: return launch_f<VelF>(P, st, cfg, stream);
        case BQ_F_DIS: return launch_f<DisF>(P, st, cfg, stream);
        case BQ_F_PROFILE: return launch_f<ProfileF>(P, st, cfg, stream);
        default: {
            // codes >= BQ_F_CUSTOM: the kernel instantiated for the user's expression lives in its
            // own small library, found through the job's integrand code
            const bq_custom_launch_fn fn = P.integrand >= BQ_F_CUSTOM ? custom_launch(P.integrand) : nullptr;
            if (!fn) return cudaErrorInvalidValue;
            return static_cast<cudaError_t>(fn(&P, &st, &cfg, stream));
        }
    }
}

}  // namespace bq
