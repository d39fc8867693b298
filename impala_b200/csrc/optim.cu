// K6: fused global-norm gradient clip + Adam step over a flat fp32 parameter bucket.
// Replaces torch.nn.utils.clip_grad_norm_ + optimizer.step() at reference
// agents/impala/learning.py:172-176 and agents/ppo/learning.py:136-140 (Adam built at
// agents/impala/builder.py:43-44: lr 1e-4, eps 1e-5, no weight decay, no amsgrad).
//
// One cooperative launch, one pass over HBM: 16 B read + 16 B written per parameter
// (p, g, m, v in; p, clipped g, m, v out).  Phase 1: every CTA sums the squares of its
// contiguous chunk of the gradient in a fixed order and publishes the partial as one 64-bit
// word (generation tag << 32 | float bits).  Every CTA then collects all words (grid barrier
// and broadcast in one step; cooperative launch guarantees co-residency), adds them in the same
// fixed order -- so all CTAs hold the same bit-identical norm -- and phase 2 re-reads its
// chunk of the gradient (L2 hit), scales it and applies Adam in torch's operation order.
// The generation counter lives in the workspace and is bumped by CTA 0 after the barrier:
// no reset pass, no host state.
#include "common.cuh"

namespace b200rl {

int env_int(const char *name, int dflt);

constexpr int kOptThreads = 256;

struct AdamParams {
    float *p, *g, *m, *v;
    long long n, chunk;
    float max_norm, one_minus_beta1, beta2, one_minus_beta2, step_size, inv_bc2_sqrt, eps;
    float *total_norm;
    unsigned int *gen;
    unsigned long long *words;
};

__device__ __forceinline__ float adam_update(float &p, float &g, float &m, float &v, float coef, const AdamParams &a) {
    g = g * coef;
    m = m + a.one_minus_beta1 * (g - m);                  // exp_avg.lerp_(grad, 1 - beta1)
    v = __fmaf_rn(a.one_minus_beta2 * g, g, v * a.beta2);  // exp_avg_sq.mul_(beta2).addcmul_(g, g, value=1-beta2)
    const float denom = sqrtf(v) * a.inv_bc2_sqrt + a.eps;
    p = p - a.step_size * (m / denom);                    // param.addcdiv_(exp_avg, denom, value=-step_size)
    return p;
}

// VEC: 16-byte aligned buffers.  RESIDENT (VEC only): the chunk is at most kResident float4 per thread, so p, g, m and v
// are all fetched BEFORE the grid barrier and stay in registers across it -- after the barrier there is only
// arithmetic and stores (the small-bucket case: 630,352 parameters are 154 such chunks).
constexpr int kResident = 4;

template <bool VEC, bool RESIDENT>
__global__ void __launch_bounds__(kOptThreads) clip_adam_kernel(const AdamParams a) {
    __shared__ float red[kOptThreads];
    __shared__ float s_total;
    const long long lo = (long long)blockIdx.x * a.chunk;
    const long long hi = min(a.n, lo + a.chunk);
    __shared__ unsigned int s_tag;
    if (threadIdx.x == 0) {
        unsigned int t = *reinterpret_cast<volatile unsigned int *>(a.gen) + 1u;
        s_tag = t ? t : 1u;  // 0 is the "never written" tag of a zeroed workspace
    }
    // ---- phase 1: partial sum of squares, fixed order ----
    float ss[1] = {0.f};
    float4 rp[kResident], rg[kResident], rm[kResident], rv[kResident];
    if (RESIDENT) {
        const float4 *p4 = reinterpret_cast<const float4 *>(a.p), *g4 = reinterpret_cast<const float4 *>(a.g);
        const float4 *m4 = reinterpret_cast<const float4 *>(a.m), *v4 = reinterpret_cast<const float4 *>(a.v);
#pragma unroll
        for (int k = 0; k < kResident; ++k) {
            const long long i = lo / 4 + threadIdx.x + (long long)k * kOptThreads;
            if (i < hi / 4) { rg[k] = g4[i]; rp[k] = p4[i]; rm[k] = m4[i]; rv[k] = v4[i]; }
            else rg[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int k = 0; k < kResident; ++k)  // same per-thread order as the streaming path
            ss[0] += (rg[k].x * rg[k].x + rg[k].y * rg[k].y) + (rg[k].z * rg[k].z + rg[k].w * rg[k].w);
        for (long long i = (hi / 4) * 4 + threadIdx.x; i < hi; i += kOptThreads) ss[0] += a.g[i] * a.g[i];  // tail of n
    } else if (VEC) {
        const float4 *g4 = reinterpret_cast<const float4 *>(a.g);
        for (long long i = lo / 4 + threadIdx.x; i < hi / 4; i += kOptThreads) {
            const float4 x = g4[i];
            ss[0] += (x.x * x.x + x.y * x.y) + (x.z * x.z + x.w * x.w);
        }
        for (long long i = (hi / 4) * 4 + threadIdx.x; i < hi; i += kOptThreads) ss[0] += a.g[i] * a.g[i];  // tail of n
    } else {
        for (long long i = lo + threadIdx.x; i < hi; i += kOptThreads) ss[0] += a.g[i] * a.g[i];
    }
    float tot[1];
    cta_sum<1>(ss, red, tot);
    const unsigned int tag = s_tag;  // written before the barrier inside cta_sum
    if (threadIdx.x == 0)
        st_volatile_u64(&a.words[blockIdx.x], ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(tot[0]));
    // ---- grid barrier + broadcast: collect every CTA's word ----
    float acc[1] = {0.f};
    const long long t0 = clock64();
    for (unsigned int c = threadIdx.x; c < gridDim.x; c += kOptThreads) {
        unsigned long long w;
        do {
            w = ld_volatile_u64(&a.words[c]);
        } while ((unsigned int)(w >> 32) != tag && clock64() - t0 < kSpinLimitCycles);
        acc[0] += ((unsigned int)(w >> 32) == tag) ? __uint_as_float((unsigned int)w) : __int_as_float(0x7fc00000);
    }
    __syncthreads();  // warp 0 done with `red`
    cta_sum<1>(acc, red, tot);
    if (threadIdx.x == 0) s_total = tot[0];
    __syncthreads();
    const float norm = sqrtf(s_total);
    float coef = 1.f;
    if (a.max_norm > 0.f) coef = fminf(a.max_norm / (norm + 1e-6f), 1.f);  // torch clip_grad_norm_
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *a.total_norm = norm;
        *reinterpret_cast<volatile unsigned int *>(a.gen) = tag;  // every CTA has read the old value (barrier above)
    }
    // ---- phase 2: clip + Adam ----
    if (RESIDENT) {
        float4 *p4 = reinterpret_cast<float4 *>(a.p), *g4 = reinterpret_cast<float4 *>(a.g);
        float4 *m4 = reinterpret_cast<float4 *>(a.m), *v4 = reinterpret_cast<float4 *>(a.v);
#pragma unroll
        for (int k = 0; k < kResident; ++k) {
            const long long i = lo / 4 + threadIdx.x + (long long)k * kOptThreads;
            if (i < hi / 4) {
                adam_update(rp[k].x, rg[k].x, rm[k].x, rv[k].x, coef, a);
                adam_update(rp[k].y, rg[k].y, rm[k].y, rv[k].y, coef, a);
                adam_update(rp[k].z, rg[k].z, rm[k].z, rv[k].z, coef, a);
                adam_update(rp[k].w, rg[k].w, rm[k].w, rv[k].w, coef, a);
                p4[i] = rp[k]; g4[i] = rg[k]; m4[i] = rm[k]; v4[i] = rv[k];
            }
        }
        for (long long i = (hi / 4) * 4 + threadIdx.x; i < hi; i += kOptThreads)
            adam_update(a.p[i], a.g[i], a.m[i], a.v[i], coef, a);
    } else if (VEC) {
        float4 *p4 = reinterpret_cast<float4 *>(a.p), *g4 = reinterpret_cast<float4 *>(a.g);
        float4 *m4 = reinterpret_cast<float4 *>(a.m), *v4 = reinterpret_cast<float4 *>(a.v);
        // backwards: the end of the chunk is what phase 1 read last, so it is the part of g most likely still in L2
        const long long first = lo / 4, count = hi / 4 - first;
        for (long long r = count - 1 - threadIdx.x; r >= 0; r -= kOptThreads) {
            const long long i = first + r;
            float4 p = p4[i], g = g4[i], m = m4[i], v = v4[i];
            adam_update(p.x, g.x, m.x, v.x, coef, a);
            adam_update(p.y, g.y, m.y, v.y, coef, a);
            adam_update(p.z, g.z, m.z, v.z, coef, a);
            adam_update(p.w, g.w, m.w, v.w, coef, a);
            p4[i] = p; g4[i] = g; m4[i] = m; v4[i] = v;
        }
        for (long long i = (hi / 4) * 4 + threadIdx.x; i < hi; i += kOptThreads)
            adam_update(a.p[i], a.g[i], a.m[i], a.v[i], coef, a);
    } else {
        for (long long i = lo + threadIdx.x; i < hi; i += kOptThreads)
            adam_update(a.p[i], a.g[i], a.m[i], a.v[i], coef, a);
    }
}

}  // namespace b200rl

extern "C" int32_t b200rl_clip_adam_step(float *params, float *grads, float *exp_avg, float *exp_avg_sq,
                                         int64_t n, float max_norm, double lr, double beta1, double beta2,
                                         double eps, int64_t step, float *total_norm, void *workspace,
                                         void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(params && grads && exp_avg && exp_avg_sq && total_norm && workspace, -1,
                   "clip_adam_step: null pointer argument");
    B200RL_REQUIRE(n >= 1, -2, "clip_adam_step: n=%lld", (long long)n);
    B200RL_REQUIRE(step >= 1, -2, "clip_adam_step: step must be >= 1 (got %lld)", (long long)step);
    B200RL_REQUIRE(lr >= 0 && beta1 >= 0 && beta1 < 1 && beta2 >= 0 && beta2 < 1 && eps >= 0, -2,
                   "clip_adam_step: invalid hyper-parameters");
    static int max_grid_dev[kMaxDevices];  // co-resident CTAs of this kernel, per device
    static int max_grid_res_dev[kMaxDevices];  // same for the register-resident instantiation
    int dev = 0;
    cudaGetDevice(&dev);
    int &max_grid = max_grid_dev[(dev >= 0 && dev < kMaxDevices) ? dev : 0];
    if (max_grid == 0) {
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, clip_adam_kernel<true, false>, kOptThreads, 0);
        if (sms <= 0 || per_sm <= 0) {
            (void)cudaGetLastError();
            set_error("clip_adam_step: occupancy query failed (no CUDA device?)");
            return (int32_t)cudaErrorUnknown;
        }
        max_grid = sms * (per_sm > 4 ? 4 : per_sm);
        if (max_grid > kMaxOptCtas) max_grid = kMaxOptCtas;
        int per_sm_res = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_res, clip_adam_kernel<true, true>, kOptThreads, 0);
        max_grid_res_dev[(dev >= 0 && dev < kMaxDevices) ? dev : 0] = sms * (per_sm_res > 0 ? per_sm_res : 0);
    }
    AdamParams a;
    a.p = params; a.g = grads; a.m = exp_avg; a.v = exp_avg_sq; a.n = n;
    const long long per_cta = env_int("B200RL_K6_ELEMS", 16 * kOptThreads);  // elements per CTA before the grid is capped (swept: scripts/k6_sweep.py)
    long long grid = (n + per_cta - 1) / per_cta;
    if (grid > max_grid) grid = max_grid;
    long long chunk = (n + grid - 1) / grid;
    chunk = (chunk + 3) / 4 * 4;  // chunks start on float4 boundaries
    grid = (n + chunk - 1) / chunk;
    a.chunk = chunk;
    a.max_norm = max_norm;
    a.one_minus_beta1 = (float)(1.0 - beta1);
    a.beta2 = (float)beta2;
    a.one_minus_beta2 = (float)(1.0 - beta2);
    const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
    a.step_size = (float)(lr / bc1);
    a.inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
    a.eps = (float)eps;
    a.total_norm = total_norm;
    a.gen = reinterpret_cast<unsigned int *>(reinterpret_cast<char *>(workspace) + kOptOffset);
    a.words = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(workspace) + kOptOffset + 64);
    const bool vec = ((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads) |
                       reinterpret_cast<uintptr_t>(exp_avg) | reinterpret_cast<uintptr_t>(exp_avg_sq)) % 16) == 0;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kOptThreads);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;  // co-residency: the in-kernel grid barrier cannot deadlock
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const bool resident = vec && chunk <= (long long)kResident * 4 * kOptThreads &&
                          grid <= max_grid_res_dev[(dev >= 0 && dev < kMaxDevices) ? dev : 0] && env_int("B200RL_K6_RESIDENT", 1);
    cudaError_t e = resident ? cudaLaunchKernelEx(&cfg, clip_adam_kernel<true, true>, a)
                    : vec    ? cudaLaunchKernelEx(&cfg, clip_adam_kernel<true, false>, a)
                             : cudaLaunchKernelEx(&cfg, clip_adam_kernel<false, false>, a);
    if (e != cudaSuccess) {
        set_error("clip_adam_kernel launch: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    return check_launch("clip_adam_kernel");
}
