// N4b: quantile-regression (Huber) loss, forward and backward in one launch.
// Replaces reference models/quantile_layers.py:83-93 quantile_regression_loss(dist_src, tau_src,
// dist_target, huber_param) as used by DistributionalApex (agents/dqn/learning.py:209):
//   delta_ij = target_j - src_i;  w_ij = |tau_i - 1[delta_ij < 0]| (detached)
//   h(d) = |d| (kappa = 0)  or  0.5 q^2 + kappa (|d| - q), q = min(|d|, kappa)
//   loss_b = mean_j sum_i w_ij h(delta_ij);  d loss_b / d src_i = -(1/M) sum_j w_ij h'(delta_ij)
// One warp per sample b: lanes stride over the N source quantiles, every lane walks the M targets
// (staged in shared memory, broadcast reads).  Deterministic (fixed-order sums).
#include "common.cuh"

namespace b200rl {

constexpr int kQhWarps = 8;

__global__ void __launch_bounds__(kQhWarps * 32) quantile_huber_kernel(
    const float *__restrict__ src, const float *__restrict__ tau, const float *__restrict__ tgt, int B, int N,
    int M, float kappa, float *__restrict__ loss, float *__restrict__ d_src) {
    extern __shared__ float s_tgt[];  // [kQhWarps][M]
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int b = blockIdx.x * kQhWarps + w;
    if (b >= B) return;
    float *t = s_tgt + (size_t)w * M;
    for (int j = lane; j < M; j += 32) t[j] = __ldg(tgt + (size_t)b * M + j);
    __syncwarp();
    const float inv_m = 1.f / (float)M;
    float acc = 0.f;
    for (int i = lane; i < N; i += 32) {
        const float x = __ldg(src + (size_t)b * N + i), ta = __ldg(tau + (size_t)b * N + i);
        float l = 0.f, g = 0.f;
        for (int j = 0; j < M; ++j) {
            const float d = t[j] - x;
            const float wgt = fabsf(ta - (d < 0.f ? 1.f : 0.f));
            const float ad = fabsf(d);
            float h, dh;  // h(d), h'(d)
            if (kappa == 0.f) {
                h = ad;
                dh = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
            } else {
                const float q = fminf(ad, kappa);
                h = 0.5f * q * q + kappa * (ad - q);
                dh = fminf(fmaxf(d, -kappa), kappa);
            }
            l = fmaf(wgt, h, l);
            g = fmaf(wgt, dh, g);
        }
        acc += l;
        d_src[(size_t)b * N + i] = -g * inv_m;  // d delta / d src = -1
    }
    acc = warp_sum(acc);
    if (lane == 0) loss[b] = acc * inv_m;
}

}  // namespace b200rl

extern "C" int32_t b200rl_quantile_huber_fwdbwd(const float *dist_src, const float *tau_src, const float *dist_target,
                                                int64_t B, int64_t N, int64_t M, float huber_param, float *loss,
                                                float *d_src, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(dist_src && tau_src && dist_target && loss && d_src, -1, "quantile_huber: null pointer argument");
    B200RL_REQUIRE(B >= 1 && N >= 1 && M >= 1, -2, "quantile_huber: need B,N,M >= 1 (got %lld,%lld,%lld)",
                   (long long)B, (long long)N, (long long)M);
    B200RL_REQUIRE(huber_param >= 0.f, -2, "quantile_huber: huber_param < 0");
    B200RL_REQUIRE(M <= 4096 && B < (1ll << 31) && N < (1ll << 31), -3, "quantile_huber: M=%lld > 4096 targets not supported", (long long)M);
    const size_t smem = (size_t)kQhWarps * M * 4;
    static size_t smem_cache[kMaxDevices];
    if (int32_t rc = ensure_dynamic_smem(quantile_huber_kernel, smem, smem_cache, "quantile_huber")) return rc;
    quantile_huber_kernel<<<(unsigned)((B + kQhWarps - 1) / kQhWarps), kQhWarps * 32, smem, (cudaStream_t)stream>>>(
        dist_src, tau_src, dist_target, (int)B, (int)N, (int)M, huber_param, loss, d_src);
    return check_launch("quantile_huber_kernel");
}
