// N3b: the prioritized-replay-weighted TD loss of the Ape-X learner, forward and backward in one launch.
// Replaces reference agents/dqn/learning.py:161-170 (after q = model(obs)) and the priorities of :172:
//   q_i   = q[i, a_i]                                      (:161-162)
//   w_i   = p_i^(-beta) / max_j p_j^(-beta)                (:164-166; importance-sampling weights of the PER sampler's
//                                                           probabilities, constants for autograd)
//   err_i = target_i - q_i                                 (:167)
//   loss  = 0.5 * mean_i(err_i^2 * w_i)                    (:168)
//   d loss / d q[i, a] = -err_i * w_i / N  at a = a_i, 0 elsewhere   (what loss.backward() hands to the Q network, :169)
//   priorities_i = |err_i|                                 (:172, fed back to ReplayBuffer.update = N3 sumtree_update)
// and the four means the learner logs (:188-191).
//
// No grid-wide max: w's maximum sits at the smallest probability (beta >= 0; the largest for beta < 0), and a minimum is
// exact, so every CTA scans all N probabilities itself (N is a replay batch: 512 in conf/agent/apex.yaml; 4 B each from
// L2) and all of them arrive at the same normaliser -- one launch, no atomics, deterministic.  The sums go through the
// generation-tagged grid reduction of the other fused losses.  dq rows are written a warp at a time as one contiguous
// run of 32 * A floats.
#include "common.cuh"

namespace b200rl {

int env_int(const char *name, int dflt);  // api.cu

constexpr int kPerThreads = 256;
constexpr int kPerMaxCtas = 296;  // 2 per SM: the redundant scan costs N * 4 B per CTA

struct PerParams {
    const float *q, *target, *prob;
    const long long *actions;
    float *dq, *priorities, *weights, *scalars;
    void *workspace;
    long long N;
    int A;
    float beta, inv_n;
};

__global__ void __launch_bounds__(kPerThreads) per_td_loss_kernel(const PerParams p) {
    __shared__ float red[4 * kPerThreads];
    __shared__ float s_ext[kPerThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_launch_dependents();
    pdl_wait();  // q comes from the Q network's last kernel, the probabilities from the sampler

    // ---- the normaliser: max_j p_j^(-beta) = (min_j p_j)^(-beta) for beta >= 0 ------------------------------------
    const bool want_min = p.beta >= 0.f;
    float ext = want_min ? INFINITY : -INFINITY;
    bool has_nan = false;
    auto take = [&](float v) {
        has_nan |= (v != v);
        ext = want_min ? fminf(ext, v) : fmaxf(ext, v);
    };
    const long long n4 = (reinterpret_cast<uintptr_t>(p.prob) & 15u) == 0 ? p.N / 4 : 0;   // 16-byte loads, 16 in flight
    const float4 *prob4 = reinterpret_cast<const float4 *>(p.prob);
#pragma unroll 16
    for (long long i = tid; i < n4; i += kPerThreads) {
        const float4 v = __ldg(prob4 + i);
        take(v.x); take(v.y); take(v.z); take(v.w);
    }
    for (long long i = n4 * 4 + tid; i < p.N; i += kPerThreads) take(__ldg(p.prob + i));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float other = __shfl_xor_sync(0xffffffffu, ext, o);
        ext = want_min ? fminf(ext, other) : fmaxf(ext, other);
    }
    has_nan = __any_sync(0xffffffffu, has_nan);
    if (lane == 0) s_ext[warp] = has_nan ? NAN : ext;
    __syncthreads();
    ext = s_ext[0];
    bool any_nan = ext != ext;
    for (int w = 1; w < kPerThreads / 32; ++w) {
        const float other = s_ext[w];
        any_nan |= (other != other);
        ext = want_min ? fminf(ext, other) : fmaxf(ext, other);
    }
    // torch's max() propagates NaN: a NaN probability makes every weight NaN there, and here
    const float wmax = any_nan ? NAN : powf(ext, -p.beta);

    // ---- this CTA's rows: a warp takes 32 consecutive rows at a time -------------------------------------------------
    float sums[4] = {0.f, 0.f, 0.f, 0.f};  // err^2 w, q, target, |err|
    const int A = p.A;
    const long long warps_total = (long long)gridDim.x * (kPerThreads / 32);
    for (long long r0 = ((long long)blockIdx.x * (kPerThreads / 32) + warp) * 32; r0 < p.N; r0 += warps_total * 32) {
        const long long i = r0 + lane;
        const bool valid = i < p.N;
        long long a = 0;
        float g = 0.f;
        if (valid) {
            a = p.actions[i];
            a = a < 0 ? 0 : (a >= A ? A - 1 : a);               // clamp like torch would raise: never read out of the row
            const float qi = __ldg(p.q + i * A + a), tg = __ldg(p.target + i);
            const float w = powf(__ldg(p.prob + i), -p.beta) / wmax;
            const float err = tg - qi;
            sums[0] += err * err * w;
            sums[1] += qi;
            sums[2] += tg;
            sums[3] += fabsf(err);
            g = -err * w * p.inv_n;
            p.priorities[i] = fabsf(err);
            if (p.weights) p.weights[i] = w;
        }
        // dq[r0 .. r0+32) as one run of rows * A floats: element e belongs to row e / A, column e % A
        const int rows = (int)min(32ll, p.N - r0);
        float *out = p.dq + r0 * A;
        const int ai = (int)a;
        for (int e0 = 0; e0 < rows * A; e0 += 32) {
            const int e = e0 + lane;
            const int row = min(e / A, 31), col = e - (e / A) * A;   // every lane takes part in the shuffles
            const float gr = __shfl_sync(0xffffffffu, g, row);
            const int ar = __shfl_sync(0xffffffffu, ai, row);
            if (e < rows * A) out[e] = (col == ar) ? gr : 0.f;
        }
    }
    const float inv_n = p.inv_n;
    float *scalars = p.scalars;
    grid_reduce<4>(sums, red, p.workspace, [=](const float(&tot)[4]) {
        scalars[0] = 0.5f * tot[0] * inv_n;  // loss, :168
        scalars[1] = tot[1] * inv_n;         // train_step/q, :188
        scalars[2] = tot[2] * inv_n;         // train_step/target, :189
        scalars[3] = tot[3] * inv_n;         // train_step/priorities, :190
        scalars[4] = wmax;                   // the normaliser (diagnostic)
        scalars[5] = scalars[6] = scalars[7] = 0.f;
    }, []() {});
}

}  // namespace b200rl

extern "C" int32_t b200rl_per_td_loss_fwdbwd(const float *q, const int64_t *actions, const float *target,
                                             const float *probabilities, int64_t N, int64_t A,
                                             float importance_sampling_exponent, float *dq, float *priorities,
                                             float *weights, float *scalars, void *workspace, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(q && actions && target && probabilities && dq && priorities && scalars && workspace, -1,
                   "per_td_loss: null pointer argument");
    B200RL_REQUIRE(N >= 1 && A >= 1, -2, "per_td_loss: need N>=1, A>=1 (got N=%lld A=%lld)", (long long)N, (long long)A);
    B200RL_REQUIRE(A <= (1 << 20) && N < (1ll << 40), -3, "per_td_loss: N=%lld A=%lld too large", (long long)N, (long long)A);
    PerParams p;
    p.q = q; p.target = target; p.prob = probabilities;
    p.actions = reinterpret_cast<const long long *>(actions);
    p.dq = dq; p.priorities = priorities; p.weights = weights; p.scalars = scalars; p.workspace = workspace;
    p.N = N; p.A = (int)A;
    p.beta = importance_sampling_exponent;
    p.inv_n = (float)(1.0 / (double)N);
    const long long rows_per_cta = (kPerThreads / 32) * 32;
    const unsigned grid = (unsigned)std::min<long long>((N + rows_per_cta - 1) / rows_per_cta, kPerMaxCtas);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kPerThreads);
    cfg.stream = reinterpret_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = env_int("B200RL_PDL", 1) ? 1 : 0;
    (void)cudaLaunchKernelEx(&cfg, per_td_loss_kernel, p);
    return check_launch("per_td_loss_kernel");
}
