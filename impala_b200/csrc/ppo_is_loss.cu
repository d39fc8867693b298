// K7: PPO clipped surrogate on V-trace advantages ("ppo_loss_is"), forward and backward in ONE launch.
// Restates the intent of reference losses.py:62-89 (not executable as committed: functorch is not imported,
// batched_vtrace is undefined in that module): V-trace over all T steps with a bootstrap value per rollout,
// the rolled not-done mask, the PPO clip on the detached masked advantage, td = 0.5 mean (mask err)^2.
//
// Same tiling as K1 (impala_loss.cu): a CTA owns C rollouts x all T steps of a batch-major [B,T,A] batch, one
// thread per sample, logits staged by one TMA bulk copy per tensor, d loss/d logits leaves by one bulk store.
// 201 B/sample at A=15: logits 60 + pi_ref 60 + a 8 + r 4 + done 1 + v 4 in, dlogits 60 + dv 4 out.
#include "softmax.cuh"

namespace b200rl {

struct PpoIsParams {
    const float *logits, *pi_ref, *rewards, *values, *v_boot;
    const unsigned char *done;
    const long long *actions;
    float *dlogits, *dvalues, *scalars;
    void *workspace;
    int B, T, A, C;
    int bulk_full, bulk_last;
    unsigned int inv_div;  // ceil(2^32 / T)
    float gamma, lambda_, one_minus_lambda, clip_rho, clip_pg_rho, clip_coeff, entropy_cost, inv_n;
    uint32_t flags;
};

template <int AT, bool EXACT>
__global__ void __launch_bounds__(AT <= 16 ? 1024 : 256) ppo_is_loss_kernel(const PpoIsParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int T = p.T, A = EXACT ? AT : p.A, C = p.C, NS = C * T;
    float *sx = reinterpret_cast<float *>(smem_raw);         // [NS*A] logits -> dlogits
    float *sy = sx + NS * A;                                 // [NS*A] behaviour logits
    float2 *s_tg = reinterpret_cast<float2 *>(sy + NS * A);  // [NS] (clipped TD error, discount * c)
    float *s_e = reinterpret_cast<float *>(s_tg + NS);       // [NS] V-trace errors e_t
    float *red = s_e + NS;                                   // [5*blockDim]
    uint64_t *bar = reinterpret_cast<uint64_t *>(red + 5 * blockDim.x);

    pdl_launch_dependents();
    const bool late_wait = p.flags & B200RL_INPUTS_STABLE;
    if (!late_wait) pdl_wait();

    const int b0 = blockIdx.x * C;
    const int cols = min(C, p.B - b0);
    const int s = threadIdx.x;
    const int c = T == 1 ? s : (int)__umulhi((unsigned)s, p.inv_div), t = s - c * T;  // ceil(2^32 / 1) does not fit
    const bool valid = s < NS && c < cols;
    const int n = (b0 + c) * T + t;

    // ---- stage the tile: one contiguous run of cols*T*A floats per tensor ----------------
    const int elems = cols * T * A;
    const size_t g0 = (size_t)b0 * T * A;
    const uint32_t bytes = (uint32_t)elems * 4u;
    const bool bulk_ok = (cols == C) ? p.bulk_full : p.bulk_last;
    if (bulk_ok) {
        if (threadIdx.x == 0) {
            mbar_init(bar, 1);
            fence_mbar_init();
            mbar_expect_tx(bar, 2u * bytes);
            bulk_g2s(sx, p.logits + g0, bytes, bar);
            bulk_g2s(sy, p.pi_ref + g0, bytes, bar);
        }
    } else {
        for (int i = threadIdx.x; i < elems; i += blockDim.x) {
            sx[i] = ld_stream_f32(p.logits + g0 + i);
            sy[i] = ld_stream_f32(p.pi_ref + g0 + i);
        }
    }
    long long act = 0;
    float rew = 0.f, disc = 0.f, mask = 0.f, v = 0.f, v_next = 0.f;
    const bool last = t == T - 1;
    if (valid) {
        act = __ldg(p.actions + n);
        rew = __ldg(p.rewards + n);
        v = __ldg(p.values + n);
        v_next = last ? __ldg(p.v_boot + b0 + c) : __ldg(p.values + n + 1);  // v_t = cat(v_tm1[:,1:], v_bootstrap), :70-72
        disc = (1.f - (float)(__ldg(p.done + n) != 0)) * p.gamma;             // :65
        // mask = (1 - d).roll(shifts=1, dims=1): step t sees the done flag of step t-1, step 0 that of step T-1 (:64)
        mask = 1.f - (float)(__ldg(p.done + (t == 0 ? n + T - 1 : n - 1)) != 0);
    }
    __syncthreads();
    if (bulk_ok) mbar_wait(bar, 0);

    // ---- per-sample statistics, TD error, trace coefficient (rho detached, :73) ----------
    SampleStats<AT> st;
    if (valid) {
        sample_stats<AT, EXACT>(st, sx + s * A, sy + s * A, A, act);
        s_tg[s] = make_float2(fminf(st.rho, p.clip_rho) * (rew + disc * v_next - v),
                              disc * (fminf(st.rho, 1.0f) * p.lambda_));
    }
    __syncthreads();

    // ---- e_k = td_k + (discount_k c_k) e_{k+1}, k = T-1 .. 0, one register chain per rollout -----
    if (s < cols) {
        const float2 *__restrict__ qp = s_tg + s * T + (T - 1);
        float *__restrict__ ep = s_e + s * T + (T - 1);
        float e = 0.f;
        int k = T - 1;
        for (; k >= 3; k -= 4, qp -= 4, ep -= 4) {
            const float2 q0 = qp[0], q1 = qp[-1], q2 = qp[-2], q3 = qp[-3];
            e = q0.x + q0.y * e; ep[0] = e;
            e = q1.x + q1.y * e; ep[-1] = e;
            e = q2.x + q2.y * e; ep[-2] = e;
            e = q3.x + q3.y * e; ep[-3] = e;
        }
        for (; k >= 0; --k, --qp, --ep) {
            const float2 q = *qp;
            e = q.x + q.y * e;
            *ep = e;
        }
    }
    __syncthreads();

    // ---- losses and gradients ---------------------------------------------------------------
    float sums[5] = {0.f, 0.f, 0.f, 0.f, 0.f};  // pg, (mask err)^2, entropy, kl, ratio
    if (valid) {
        const float e = s_e[s];
        const float err = (e + v) - v;  // errors = stop_gradient(e + v_tm1) - v_tm1
        float qb = v_next;
        if (!last) {
            const float tgt1 = ((s_e[s + 1] + v_next) - v_next) + v_next;
            // two rounded products and a rounded sum, as the reference's tensor expression evaluates it: never contracted
            // into an FMA (which of the two products the compiler fuses differed between kernels -- tests/test_gpu_fuzz.py)
            qb = __fadd_rn(__fmul_rn(p.lambda_, tgt1), __fmul_rn(p.one_minus_lambda, v_next));
        }
        const float q = rew + disc * qb;
        const float adv = (fminf(st.rho, p.clip_pg_rho) * (q - v)) * mask;  // (adv * mask).detach(), :75
        const float ratio = st.rho;
        const float lo = 1.f - p.clip_coeff, hi = 1.f + p.clip_coeff;
        const float pg1 = -(adv * ratio);                       // :76
        const float pg2 = -fminf(fmaxf(ratio, lo), hi) * adv;   // :77
        const float me = mask * err;
        sums[0] = fmaxf(pg1, pg2);                              // :78
        sums[1] = me * me;                                      // :74
        sums[2] = st.H;
        sums[3] = st.kl;
        sums[4] = ratio;
        // same tie rule as K3 (ppo_loss.cu): inside the clip range both branches carry -adv
        const bool inside = ratio >= lo && ratio <= hi;
        const float g_ratio = (inside || pg1 > pg2) ? -adv : 0.f;
        p.dvalues[n] = -(mask * me) * p.inv_n;                  // d(0.5 (mask err)^2)/dv; v_t and adv carry no gradient
        write_dlogits<AT, EXACT>(st, sx + s * A, A, g_ratio * ratio * p.inv_n, p.entropy_cost * p.inv_n);
    }
    if (bulk_ok) fence_proxy_async_smem();
    __syncthreads();
    if (bulk_ok) {
        if (threadIdx.x == 0) {
            bulk_s2g(p.dlogits + g0, sx, bytes);
            bulk_commit();
        }
    } else {
        for (int i = threadIdx.x; i < elems; i += blockDim.x) p.dlogits[g0 + i] = sx[i];
    }
    const float inv_n = p.inv_n, ec = p.entropy_cost;
    float *scalars = p.scalars;
    grid_reduce<5>(
        sums, red, p.workspace,
        [=](const float(&tot)[5]) {
            const float pg = tot[0] * inv_n, td = 0.5f * tot[1] * inv_n, ent = tot[2] * inv_n;
            scalars[B200RL_S_LOSS] = pg + td - ec * ent;  // :81
            scalars[B200RL_S_ENTROPY] = ent;
            scalars[B200RL_S_TD] = td;
            scalars[B200RL_S_PG] = pg;
            scalars[B200RL_S_KL] = tot[3] * inv_n;  // unclamped: clamp_min(0.) (:79) belongs after the cross-rank sum
            scalars[B200RL_S_RATIO] = tot[4] * inv_n;
            scalars[6] = 0.f;
            scalars[7] = 0.f;
        },
        [=]() { if (late_wait) pdl_wait(); });
    if (bulk_ok && threadIdx.x == 0) bulk_wait_read0();
}

int env_int(const char *name, int dflt);

template <int AT, bool EXACT>
static int32_t launch_ppo_is(const PpoIsParams &p, int threads, size_t smem, cudaStream_t stream) {
    static size_t smem_cache[kMaxDevices];
    if (int32_t rc = ensure_dynamic_smem(ppo_is_loss_kernel<AT, EXACT>, smem, smem_cache, "ppo_is_loss")) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((p.B + p.C - 1) / p.C));
    cfg.blockDim = dim3((unsigned)threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = env_int("B200RL_PDL", 1) ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&cfg, ppo_is_loss_kernel<AT, EXACT>, p);
    if (e != cudaSuccess) {
        set_error("ppo_is_loss_kernel launch: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    return check_launch("ppo_is_loss_kernel");
}

}  // namespace b200rl

extern "C" int32_t b200rl_ppo_is_loss_fwdbwd(
    const float *logits, const float *pi_ref, const int64_t *actions, const float *rewards,
    const uint8_t *done, const float *values, const float *v_boot, int64_t B, int64_t T, int64_t A,
    int64_t n_global_b, float gamma, double lambda_, float clip_rho, float clip_pg_rho, float clip_coeff,
    float entropy_cost, uint32_t flags, float *dlogits, float *dvalues, float *scalars, void *workspace,
    void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(logits && pi_ref && actions && rewards && done && values && v_boot && dlogits && dvalues &&
                       scalars && workspace, -1, "ppo_is_loss: null pointer argument");
    B200RL_REQUIRE(B >= 1 && T >= 1 && A >= 1, -2, "ppo_is_loss: need B>=1, T>=1, A>=1 (got B=%lld T=%lld A=%lld)",
                   (long long)B, (long long)T, (long long)A);
    B200RL_REQUIRE(A <= 32, -3, "ppo_is_loss: A=%lld > 32 actions not supported", (long long)A);
    const int max_threads = A <= 16 ? 640 : 256;
    B200RL_REQUIRE(T <= max_threads, -3, "ppo_is_loss: T=%lld > %d steps not supported at A=%lld", (long long)T,
                   max_threads, (long long)A);
    B200RL_REQUIRE(B * T < (1ll << 31), -3, "ppo_is_loss: B*T too large");
    B200RL_REQUIRE(n_global_b >= B, -2, "ppo_is_loss: n_global_b < B");

    PpoIsParams p;
    p.logits = logits; p.pi_ref = pi_ref; p.rewards = rewards; p.values = values; p.v_boot = v_boot;
    p.done = done; p.actions = reinterpret_cast<const long long *>(actions);
    p.dlogits = dlogits; p.dvalues = dvalues; p.scalars = scalars; p.workspace = workspace;
    p.B = (int)B; p.T = (int)T; p.A = (int)A;
    p.gamma = gamma; p.lambda_ = (float)lambda_; p.one_minus_lambda = (float)(1.0 - lambda_);
    p.clip_rho = clip_rho; p.clip_pg_rho = clip_pg_rho; p.clip_coeff = clip_coeff; p.entropy_cost = entropy_cost;
    p.inv_n = (float)(1.0 / ((double)n_global_b * (double)T));
    p.flags = flags;

    int C = env_int("B200RL_K1_SAMPLES", 160) / (int)T;
    if (C < 1) C = 1;
    if (C > B) C = (int)B;
    while (C > 1 && C * T > max_threads) --C;
    auto threads_for = [&](int c) { return ((c * (int)T + 31) / 32) * 32; };
    auto smem_for = [&](int c) { return (size_t)c * T * (2 * A + 3) * 4 + 5 * (size_t)threads_for(c) * 4 + 16; };
    while (C > 1 && smem_for(C) > 200 * 1024) --C;
    B200RL_REQUIRE(smem_for(C) <= 226 * 1024, -3, "ppo_is_loss: T*A too large for one CTA tile");
    if ((B + C - 1) / C > kMaxPartialCtas) {  // wide batches of long rollouts: wider CTA tiles so the grid fits the workspace's partial sums
        const int need = (int)((B + kMaxPartialCtas - 1) / kMaxPartialCtas);
        B200RL_REQUIRE((long long)need * T <= max_threads && smem_for(need) <= 226 * 1024, -3,
                       "ppo_is_loss: B=%lld columns of T=%lld need %d columns per CTA to fit the %d partial sums of the workspace, "
                       "which one CTA cannot hold -- split the batch", (long long)B, (long long)T, need, kMaxPartialCtas);
        C = need;
    }
    p.C = C;
    const bool ptrs = ((reinterpret_cast<uintptr_t>(logits) | reinterpret_cast<uintptr_t>(pi_ref) |
                        reinterpret_cast<uintptr_t>(dlogits)) % 16) == 0;
    auto ok = [&](long long cols) {
        return (ptrs && (cols * T * A * 4) % 16 == 0 && ((long long)C * T * A * 4) % 16 == 0) ? 1 : 0;
    };
    p.bulk_full = ok(C);
    p.bulk_last = (B % C) ? ok(B % C) : p.bulk_full;
    p.inv_div = (unsigned int)(((1ull << 32) + (unsigned)T - 1) / (unsigned)T);
    const int threads = threads_for(C);
    const size_t smem = smem_for(C);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    switch (A) {
        case 6: return launch_ppo_is<6, true>(p, threads, smem, s);
        case 15: return launch_ppo_is<15, true>(p, threads, smem, s);
        case 18: return launch_ppo_is<18, true>(p, threads, smem, s);
        default: break;
    }
    if (A <= 8) return launch_ppo_is<8, false>(p, threads, smem, s);
    if (A <= 16) return launch_ppo_is<16, false>(p, threads, smem, s);
    return launch_ppo_is<32, false>(p, threads, smem, s);
}
