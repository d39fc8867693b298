// K3: fused PPO clipped-surrogate loss, forward and backward in one launch.
// Replaces reference losses.py:134-155 (after model(s)) and the loss part of
// loss.backward() at agents/ppo/learning.py:134.
//
// Flat minibatch [N,A]: a CTA stages a tile of kTile samples' logits and behaviour
// logits in shared memory with one TMA bulk copy per tensor, one thread per sample
// computes everything in registers, d loss/d logits is written over the staged tile
// and leaves by one bulk store.  200 B/sample at A=15.
#include "ppo_loss.cuh"
#include "softmax.cuh"

namespace b200rl {

constexpr int kPpoTile = 256;

template <int AT, bool EXACT>
__global__ void __launch_bounds__(kPpoTile) ppo_loss_kernel(const PpoParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int A = EXACT ? AT : p.A;
    float *sx = reinterpret_cast<float *>(smem_raw);
    float *sy = sx + (size_t)kPpoTile * A;
    float *red = sy + (size_t)kPpoTile * A;  // [6*kPpoTile]
    uint64_t *bar = reinterpret_cast<uint64_t *>(red + 6 * kPpoTile);

    pdl_launch_dependents();
    const bool late_wait = p.flags & B200RL_INPUTS_STABLE;
    if (!late_wait) pdl_wait();  // inputs may come from the preceding kernel


    const long long n0 = (long long)blockIdx.x * kPpoTile;
    const int count = (int)min((long long)kPpoTile, p.N - n0);
    const int s = threadIdx.x;
    const bool valid = s < count;
    const size_t g0 = (size_t)n0 * A, elems = (size_t)count * A;
    const uint32_t bytes = (uint32_t)(elems * 4);
    const bool bulk_ok = (bytes % 16 == 0) && ((g0 * 4) % 16 == 0) &&
                         ((reinterpret_cast<uintptr_t>(p.logits) | reinterpret_cast<uintptr_t>(p.pi_ref) |
                           reinterpret_cast<uintptr_t>(p.dlogits)) % 16 == 0);
    if (bulk_ok) {
        if (threadIdx.x == 0) {
            mbar_init(bar, 1);
            fence_mbar_init();
            mbar_expect_tx(bar, 2u * bytes);
            bulk_g2s(sx, p.logits + g0, bytes, bar);
            bulk_g2s(sy, p.pi_ref + g0, bytes, bar);
        }
    } else {
        for (size_t i = threadIdx.x; i < elems; i += blockDim.x) {
            sx[i] = ld_stream_f32(p.logits + g0 + i);
            sy[i] = ld_stream_f32(p.pi_ref + g0 + i);
        }
    }
    long long act = 0;
    float vt = 0.f, v = 0.f;
    if (valid) {
        act = p.actions[n0 + s];
        vt = p.v_target[n0 + s];
        v = p.v[n0 + s];
    }
    __syncthreads();
    if (bulk_ok) mbar_wait(bar, 0);

    float sums[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // pg, td, entropy, kl, ratio, target
    if (valid) {
        SampleStats<AT> st;
        sample_stats<AT, EXACT>(st, sx + s * A, sy + s * A, A, act);
        const float ratio = st.rho;
        const float adv = vt - v;                                   // losses.py:137
        const float lo = 1.f - p.clip_coeff, hi = 1.f + p.clip_coeff;
        const float pg1 = -(adv * ratio);                           // :140
        const float pg2 = -fminf(fmaxf(ratio, lo), hi) * adv;       // :141
        sums[0] = fmaxf(pg1, pg2);                                  // :142
        sums[1] = adv * adv;
        sums[2] = st.H;
        sums[3] = st.kl;
        sums[4] = ratio;
        sums[5] = vt;
        // torch.max splits ties evenly and clamp passes gradient on [lo,hi]: inside the
        // range both branches carry d/d ratio = -adv (halves add up); outside only pg1 can.
        const bool inside = ratio >= lo && ratio <= hi;
        const float g_ratio = (inside || pg1 > pg2) ? -adv : 0.f;
        const float g_lp = g_ratio * ratio * p.inv_n;               // d ratio / d lp = ratio
        p.dv[n0 + s] = -adv * p.inv_n;                              // d(0.5 adv^2)/dv, adv detached in pg
        write_dlogits<AT, EXACT>(st, sx + s * A, A, g_lp, p.entropy_cost * p.inv_n);
    }
    if (bulk_ok) fence_proxy_async_smem();
    __syncthreads();
    if (bulk_ok) {
        if (threadIdx.x == 0) {
            bulk_s2g(p.dlogits + g0, sx, bytes);
            bulk_commit();
        }
    } else {
        for (size_t i = threadIdx.x; i < elems; i += blockDim.x) p.dlogits[g0 + i] = sx[i];
    }
    const float inv_n = p.inv_n, ec = p.entropy_cost;
    float *scalars = p.scalars;
    grid_reduce<6>(sums, red, p.workspace, [=](const float(&tot)[6]) {
        const float pg = tot[0] * inv_n, td = 0.5f * tot[1] * inv_n, ent = tot[2] * inv_n;
        scalars[B200RL_S_LOSS] = pg + td - ec * ent;  // losses.py:146
        scalars[B200RL_S_ENTROPY] = ent;
        scalars[B200RL_S_TD] = td;
        scalars[B200RL_S_PG] = pg;
        scalars[B200RL_S_KL] = tot[3] * inv_n;
        scalars[B200RL_S_RATIO] = tot[4] * inv_n;
        scalars[B200RL_S_TARGET] = tot[5] * inv_n;
        scalars[7] = 0.f;
    }, [=]() { if (late_wait) pdl_wait(); });
    if (bulk_ok && threadIdx.x == 0) bulk_wait_read0();
}

template <int AT, bool EXACT>
static int32_t launch_ppo(const PpoParams &p, size_t smem, cudaStream_t stream) {
    static size_t smem_cache[kMaxDevices];
    if (int32_t rc = ensure_dynamic_smem(ppo_loss_kernel<AT, EXACT>, smem, smem_cache, "ppo_loss")) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((p.N + kPpoTile - 1) / kPpoTile));
    cfg.blockDim = dim3(kPpoTile);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // PDL: see pdl_wait() in the kernel
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = env_int("B200RL_PDL", 1) ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&cfg, ppo_loss_kernel<AT, EXACT>, p);
    if (e != cudaSuccess) {
        set_error("ppo_loss_kernel launch: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    return check_launch("ppo_loss_kernel");
}

}  // namespace b200rl

extern "C" int32_t b200rl_ppo_loss_fwdbwd(const float *logits, const float *pi_ref, const int64_t *actions,
                                          const float *v_target, const float *v, int64_t N, int64_t A,
                                          int64_t n_global, float clip_coeff, float entropy_cost,
                                          uint32_t flags, float *dlogits, float *dv, float *scalars,
                                          void *workspace, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(logits && pi_ref && actions && v_target && v && dlogits && dv && scalars && workspace, -1,
                   "ppo_loss: null pointer argument");
    B200RL_REQUIRE(N >= 1 && A >= 1, -2, "ppo_loss: need N>=1, A>=1 (got N=%lld A=%lld)", (long long)N, (long long)A);
    B200RL_REQUIRE(A <= 32, -3, "ppo_loss: A=%lld > 32 actions not supported", (long long)A);
    B200RL_REQUIRE(n_global >= N, -2, "ppo_loss: n_global < N");
    B200RL_REQUIRE((N + kPpoTile - 1) / kPpoTile <= kMaxPartialCtas, -3,
                   "ppo_loss: N=%lld too large for the partials workspace (max %lld)", (long long)N,
                   (long long)kMaxPartialCtas * kPpoTile);
    PpoParams p;
    p.logits = logits; p.pi_ref = pi_ref; p.v_target = v_target; p.v = v;
    p.actions = reinterpret_cast<const long long *>(actions);
    p.dlogits = dlogits; p.dv = dv; p.scalars = scalars; p.workspace = workspace;
    p.N = N; p.A = (int)A; p.clip_coeff = clip_coeff; p.entropy_cost = entropy_cost;
    p.inv_n = (float)(1.0 / (double)n_global);
    p.flags = flags;
    const size_t smem = (size_t)kPpoTile * A * 8 + 6 * kPpoTile * 4 + 16;
    cudaStream_t s = (cudaStream_t)stream;
    // Two schedules (B200RL_K3_STREAM: 0 = this rule, 1 = always the tile kernel, 2 = always try the streaming kernel):
    // a launch that has the GPU to itself and is large enough to give every CTA slot a few chunks streams through the
    // persistent ring kernel; small minibatches and launches that overlap others (B200RL_INPUTS_STABLE) keep the tiles.
    const int impl = env_int("B200RL_K3_STREAM", 0);
    if (impl == 2 || (impl == 0 && !(flags & B200RL_INPUTS_STABLE))) {
        const int32_t rc = launch_ppo_stream(p, s);
        if (rc != 1) return rc;
    }
    switch (A) {
        case 4: return launch_ppo<4, true>(p, smem, s);
        case 6: return launch_ppo<6, true>(p, smem, s);
        case 15: return launch_ppo<15, true>(p, smem, s);
        case 18: return launch_ppo<18, true>(p, smem, s);
        default: break;
    }
    if (A <= 8) return launch_ppo<8, false>(p, smem, s);
    if (A <= 16) return launch_ppo<16, false>(p, smem, s);
    return launch_ppo<32, false>(p, smem, s);
}
