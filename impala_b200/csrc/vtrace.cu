// Stand-alone V-trace (drop-in for batched_vtrace, reference agents/impala/learning.py:15-26,
// 150-153; algorithm: rlax vtrace_td_error_and_advantage) and its backward.
//
// Forward (Tm <= 256): K1's tiling without the softmax -- a CTA owns C columns x all Tm steps, one thread per
// (column, step) in MEMORY order so every load and store is coalesced for either layout; the (td, discount*c)
// pairs meet in shared memory, the column's first thread runs the recursion as a register chain, and every
// thread then finishes its own step.  The first version (one thread per column walking global memory, 32 CTAs
// for B = 4096) took 48 us for the 2.5 MB of BASELINE config 2; this one is launch-bound.  Longer sequences keep
// the one-thread-per-column kernel.  (The learner's hot path does not come through here -- K1 fuses this
// recursion into the loss kernel.)
#include "common.cuh"

namespace b200rl {

__global__ void __launch_bounds__(128) vtrace_fwd_kernel(
    const float *__restrict__ v_tm1, const float *__restrict__ v_t, const float *__restrict__ r_t,
    const float *__restrict__ disc, const float *__restrict__ rho, int B, int Tm, long long sb,
    long long st, float lambda_, float oml, float clip_rho, float clip_pg_rho, float *__restrict__ adv,
    float *__restrict__ err, float *__restrict__ q) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    float e = 0.f, tgt_next = 0.f, v_tm1_next = 0.f;
    long long i = (long long)b * sb + (long long)(Tm - 1) * st;
    const long long ost = (st == 1) ? 1 : B;  // outputs are compact, same major order
    long long o = (st == 1) ? (long long)b * Tm + (Tm - 1) : (long long)(Tm - 1) * B + b;
    for (int t = Tm - 1; t >= 0; --t, i -= st, o -= ost) {
        const float rh = rho[i], vm = v_tm1[i], vt = v_t[i], r = r_t[i], d = disc[i];
        const float c = fminf(rh, 1.0f) * lambda_;
        const float td = fminf(rh, clip_rho) * (r + d * vt - vm);
        e = td + d * c * e;
        const float tgt = e + vm;
        const float er = tgt - vm;
        const float qb = (t == Tm - 1) ? vt : lambda_ * tgt_next + oml * v_tm1_next;
        const float qq = r + d * qb;
        err[o] = er;
        q[o] = qq;
        adv[o] = fminf(rh, clip_pg_rho) * (qq - vm);
        tgt_next = er + vm;
        v_tm1_next = vm;
    }
}

constexpr int kVtThreads = 256;

template <bool TM>
__global__ void __launch_bounds__(kVtThreads) vtrace_fwd_tiled_kernel(
    const float *__restrict__ v_tm1, const float *__restrict__ v_t, const float *__restrict__ r_t,
    const float *__restrict__ disc, const float *__restrict__ rho, int B, int Tm, int C, long long sb,
    long long st, float lambda_, float oml, float clip_rho, float clip_pg_rho, float *__restrict__ adv,
    float *__restrict__ err, float *__restrict__ q) {
    __shared__ float2 s_tg[kVtThreads];
    __shared__ float s_e[kVtThreads], s_vm[kVtThreads];
    const int s = threadIdx.x, b0 = blockIdx.x * C, cols = min(C, B - b0);
    int c, t;
    if (TM) { t = s / C; c = s - t * C; } else { c = s / Tm; t = s - c * Tm; }
    const bool valid = s < C * Tm && c < cols;
    const int stl = TM ? C : 1;  // local distance between steps t and t+1
    float rh = 0.f, vm = 0.f, vt = 0.f, r = 0.f, d = 0.f;
    if (valid) {
        const long long i = (long long)(b0 + c) * sb + (long long)t * st;
        rh = rho[i]; vm = v_tm1[i]; vt = v_t[i]; r = r_t[i]; d = disc[i];
        const float cc = fminf(rh, 1.0f) * lambda_;
        s_tg[s] = make_float2(fminf(rh, clip_rho) * (r + d * vt - vm), d * cc);
        s_vm[s] = vm;
    }
    __syncthreads();
    if (s < cols) {
        int k = (TM ? s : s * Tm) + (Tm - 1) * stl;
        float e = 0.f;
        for (int j = Tm - 1; j >= 0; --j, k -= stl) {
            const float2 x = s_tg[k];
            e = x.x + x.y * e;
            s_e[k] = e;
        }
    }
    __syncthreads();
    if (!valid) return;
    const float er = (s_e[s] + vm) - vm;
    float qb = vt;
    if (t < Tm - 1) {
        const float vm1 = s_vm[s + stl];
        const float tgt1 = ((s_e[s + stl] + vm1) - vm1) + vm1;
        qb = lambda_ * tgt1 + oml * vm1;
    }
    const float qq = r + d * qb;
    const long long o = TM ? (long long)t * B + b0 + c : (long long)(b0 + c) * Tm + t;  // outputs are compact
    err[o] = er;
    q[o] = qq;
    adv[o] = fminf(rh, clip_pg_rho) * (qq - vm);
}

// Elementwise backward (stop_target_gradients=True: vs targets are constants).
__global__ void __launch_bounds__(256) vtrace_bwd_kernel(
    const float *__restrict__ v_tm1, const float *__restrict__ disc, const float *__restrict__ rho,
    const float *__restrict__ q, const float *__restrict__ g_adv, const float *__restrict__ g_err,
    const float *__restrict__ g_q, int B, int Tm, long long sb, long long st, float oml,
    float clip_pg_rho, float *__restrict__ d_v_tm1, float *__restrict__ d_v_t,
    float *__restrict__ d_rho) {
    const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= (long long)B * Tm) return;
    // enumerate in memory order so accesses coalesce for either layout
    int b, t;
    if (st == 1) { b = (int)(k / Tm); t = (int)(k - (long long)b * Tm); }
    else { t = (int)(k / B); b = (int)(k - (long long)t * B); }
    const long long i = (long long)b * sb + (long long)t * st;  // strided inputs
    const long long o = k, ost = (st == 1) ? 1 : B;              // compact q / upstream / outputs
    const float rh = rho[i];
    const float cpg = fminf(rh, clip_pg_rho);
    const float ga = g_adv ? g_adv[o] : 0.f;
    const float gq = (g_q ? g_q[o] : 0.f) + ga * cpg;  // total upstream on q_t
    float dvm = -(g_err ? g_err[o] : 0.f) - ga * cpg;
    if (t > 0 && oml != 0.f) {
        const long long j = i - st, oj = o - ost;
        const float gq_prev = (g_q ? g_q[oj] : 0.f) + (g_adv ? g_adv[oj] : 0.f) * fminf(rho[j], clip_pg_rho);
        dvm += gq_prev * disc[j] * oml;
    }
    d_v_tm1[o] = dvm;
    d_v_t[o] = (t == Tm - 1) ? gq * disc[i] : 0.f;
    d_rho[o] = (rh <= clip_pg_rho) ? ga * (q[o] - v_tm1[i]) : 0.f;
}

static int32_t check_layout(const char *who, int64_t B, int64_t T, int64_t sb, int64_t st) {
    B200RL_REQUIRE(B >= 1 && T >= 1, -2, "%s: need B>=1 and T>=1 (got B=%lld T=%lld)", who, (long long)B, (long long)T);
    B200RL_REQUIRE(B * T < (1ll << 31), -3, "%s: B*T too large", who);
    B200RL_REQUIRE((st == 1 && sb >= T) || (sb == 1 && st >= B), -4,
                   "%s: strides (%lld,%lld) are neither batch-major (>=T,1) nor time-major (1,>=B)", who,
                   (long long)sb, (long long)st);
    return 0;
}

}  // namespace b200rl

extern "C" int32_t b200rl_vtrace(const float *v_tm1, const float *v_t, const float *r_t,
                                 const float *discount_t, const float *rho_tm1, int64_t B, int64_t Tm,
                                 int64_t stride_b, int64_t stride_t, double lambda_, float clip_rho,
                                 float clip_pg_rho, float *adv, float *err, float *q, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(v_tm1 && v_t && r_t && discount_t && rho_tm1 && adv && err && q, -1, "vtrace: null pointer argument");
    if (int32_t rc = check_layout("vtrace", B, Tm, stride_b, stride_t)) return rc;
    if (Tm <= kVtThreads) {
        const bool tm = !(stride_t == 1 && stride_b >= Tm);
        int C = kVtThreads / (int)Tm;
        if (C > B) C = (int)B;
        const unsigned grid = (unsigned)((B + C - 1) / C);
        if (tm)
            vtrace_fwd_tiled_kernel<true><<<grid, kVtThreads, 0, (cudaStream_t)stream>>>(
                v_tm1, v_t, r_t, discount_t, rho_tm1, (int)B, (int)Tm, C, stride_b, stride_t, (float)lambda_,
                (float)(1.0 - lambda_), clip_rho, clip_pg_rho, adv, err, q);
        else
            vtrace_fwd_tiled_kernel<false><<<grid, kVtThreads, 0, (cudaStream_t)stream>>>(
                v_tm1, v_t, r_t, discount_t, rho_tm1, (int)B, (int)Tm, C, stride_b, stride_t, (float)lambda_,
                (float)(1.0 - lambda_), clip_rho, clip_pg_rho, adv, err, q);
        return check_launch("vtrace_fwd_tiled_kernel");
    }
    const int threads = 128;
    vtrace_fwd_kernel<<<(unsigned)((B + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(
        v_tm1, v_t, r_t, discount_t, rho_tm1, (int)B, (int)Tm, stride_b, stride_t, (float)lambda_,
        (float)(1.0 - lambda_), clip_rho, clip_pg_rho, adv, err, q);
    return check_launch("vtrace_fwd_kernel");
}

extern "C" int32_t b200rl_vtrace_bwd(const float *v_tm1, const float *discount_t, const float *rho_tm1,
                                     const float *q, const float *g_adv, const float *g_err,
                                     const float *g_q, int64_t B, int64_t Tm, int64_t stride_b,
                                     int64_t stride_t, double lambda_, float clip_pg_rho, float *d_v_tm1,
                                     float *d_v_t, float *d_rho, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(v_tm1 && discount_t && rho_tm1 && q && d_v_tm1 && d_v_t && d_rho, -1, "vtrace_bwd: null pointer argument");
    if (int32_t rc = check_layout("vtrace_bwd", B, Tm, stride_b, stride_t)) return rc;
    const int threads = 256;
    const long long n = (long long)B * Tm;
    vtrace_bwd_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(
        v_tm1, discount_t, rho_tm1, q, g_adv, g_err, g_q, (int)B, (int)Tm, stride_b, stride_t,
        (float)(1.0 - lambda_), clip_pg_rho, d_v_tm1, d_v_t, d_rho);
    return check_launch("vtrace_bwd_kernel");
}
