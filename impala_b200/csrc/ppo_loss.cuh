// K3 plumbing shared by the two schedules of the fused PPO loss: ppo_loss.cu (one CTA per tile) and ppo_loss_stream.cu
// (persistent, ring-pipelined).
#pragma once
#include "common.cuh"

namespace b200rl {

struct PpoParams {
    const float *logits, *pi_ref, *v_target, *v;
    const long long *actions;
    float *dlogits, *dv, *scalars;
    void *workspace;
    long long N;
    int A;
    float clip_coeff, entropy_cost, inv_n;
    uint32_t flags;
};

int env_int(const char *name, int dflt);
// returns 1 when the shape is not taken by the streaming kernel (the caller then uses the tile kernel)
int32_t launch_ppo_stream(const PpoParams &p, cudaStream_t stream);

}  // namespace b200rl
