// K1 host-side plumbing shared by the two implementations of the fused V-trace + IMPALA loss:
//   impala_loss.cu       persistent, warp-group pipelined kernel (batch-major; the default)
//   impala_loss_tile.cu  one CTA per tile (time-major layout and shapes the persistent kernel does not take)
#pragma once
#include "common.cuh"

namespace b200rl {

// Arguments of b200rl_vtrace_impala_loss_fwdbwd after validation (include/b200rl.h).
struct ImpalaArgs {
    const float *logits, *pi_ref, *rewards, *discounts, *values;
    const long long *actions;
    float *dlogits, *dvalues, *scalars;
    void *workspace;
    int B, T, A, time_major;
    float lambda_, one_minus_lambda, clip_rho, clip_pg_rho, entropy_coeff, inv_n_pg, inv_n_all;
    uint32_t flags;
};

int env_int(const char *name, int dflt);
int32_t launch_impala_tile(const ImpalaArgs &a, cudaStream_t stream);
// returns 1 when the shape is not taken by the persistent kernel (the caller then uses the tile kernel)
int32_t launch_impala_persistent(const ImpalaArgs &a, cudaStream_t stream);

}  // namespace b200rl
