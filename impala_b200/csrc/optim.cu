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
    // CUDA-graph capture: the step-dependent values may not be launch arguments.  counters (nullable, device):
    // [0] Adam's step, [1] the cross-rank flag generation (K8).  Every CTA reads them at entry, uses value + 1 and CTA 0
    // stores value + 1 back after the grid barrier; the bias corrections are then formed in the kernel from lr / betas.
    long long *counters;
    double lr, beta1, beta2_d;
};

// ---- K8: the same kernel with the gradient ALLREDUCE fused in (PEER = true) ----------------------------------
// Every rank's flat gradient bucket is mapped into every other rank's address space (CUDA IPC over NVLink /
// NVSwitch peer access).  Instead of reading its own gradient, a CTA reads its chunk from ALL ranks' buckets
// with system-scope loads and adds them in rank order 0..W-1 -- every rank computes bit-identical sums -- then
// norm, clip and Adam proceed as in K6.  Two flag exchanges bracket the peer reads: "my gradients are final"
// before them, "I have finished reading yours" after them (the kernel does not exit before every peer has
// said so, so stream order protects the next backward pass that overwrites the bucket).  Flags carry the step
// number: monotonic, never reset.
// Two routes.  ONE-SHOT (2 ranks): every rank reads every bucket in full -- (W-1) n floats over NVLink per rank,
// one flag round trip.  TWO-SHOT (more ranks): CTA c of rank r reduces only the r-th part of chunk c
// (reduce-scatter: (W-1)/W n floats read per rank) and stores the sums into the `reduced` buffer of EVERY rank
// (all-gather: (W-1)/W n floats written), then the W CTAs "c" of the ranks exchange per-CTA flags and each reads
// its now complete local chunk.  A sum is formed once, by one rank, in rank order: bit-identical everywhere.
// NVLS (two-shot with multicast addresses of the same buffers): the part is fetched with ONE multimem.ld_reduce --
// the NVSwitch adds the W ranks' values -- and broadcast with ONE multimem.st: n/W floats each way per rank.
constexpr int kMaxPeers = 8;
constexpr long long kPeerSpinLimitCycles = 1ll << 35;  // ~18 s: ranks may be skewed by a whole learner step

struct PeerParams {
    const float *g[kMaxPeers];      // every rank's bucket: n gradients followed by n_extra floats
    unsigned int *flags[kMaxPeers];  // every rank's flag area: ready[kMaxPeers], done[kMaxPeers]
    float *red[kMaxPeers];           // two-shot: every rank's reduced-gradient buffer (red[rank] == AdamParams::g)
    unsigned int *cflags[kMaxPeers]; // two-shot: every rank's per-CTA flag area [kMaxOptCtas][kMaxPeers]
    const float *mc_g;               // NVLS: multicast address of the buckets (nullptr: unicast loads / stores)
    float *mc_red;                   // NVLS: multicast address of the reduced buffers
    float *extra_out;                // [n_extra] sums of the buckets' tails (loss / metric scalars)
    int world, rank, n_extra, two_shot;
    unsigned int gen;
};

__device__ __forceinline__ void st_release_sys_u32(unsigned int *p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys_u32(const unsigned int *p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_sys_f4(const float4 *p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_sys_f4(float4 *p, float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float4 *mc) {  // sum over the multicast group, formed in the switch
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st_f4(float4 *mc, float4 v) {  // one store, delivered to every rank
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float ld_sys_f32(const float *p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
// wait until peer `t`'s word in this rank's flag area reaches `gen`; false on timeout
__device__ __forceinline__ bool peer_wait(const unsigned int *word, unsigned int gen) {
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys_u32(word) - gen) < 0)
        if (clock64() - t0 > kPeerSpinLimitCycles) return false;
    return true;
}
template <bool PEER>
__device__ __forceinline__ float4 load_grad4(const AdamParams &a, const PeerParams &pp, long long i) {
    if constexpr (!PEER) {
        return reinterpret_cast<const float4 *>(a.g)[i];
    } else {
    if (pp.two_shot) return __ldcg(reinterpret_cast<const float4 *>(a.g) + i);  // complete after the exchange phase
    float4 v[kMaxPeers];
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r)
        if (r < pp.world) v[r] = ld_sys_f4(reinterpret_cast<const float4 *>(pp.g[r]) + i);
    float4 s = v[0];
#pragma unroll
    for (int r = 1; r < kMaxPeers; ++r)
        if (r < pp.world) { s.x += v[r].x; s.y += v[r].y; s.z += v[r].z; s.w += v[r].w; }
    return s;
    }
}
template <bool PEER>
__device__ __forceinline__ float load_grad1(const AdamParams &a, const PeerParams &pp, long long i) {
    if constexpr (!PEER) {
        return a.g[i];
    } else {
        float s = ld_sys_f32(pp.g[0] + i);
        for (int r = 1; r < pp.world; ++r) s += ld_sys_f32(pp.g[r] + i);
        return s;
    }
}

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

template <bool VEC, bool RESIDENT, bool PEER>
__global__ void __launch_bounds__(kOptThreads) clip_adam_kernel(const AdamParams a_in, const PeerParams pp_in) {
    __shared__ float red[kOptThreads];
    __shared__ float s_total;
    __shared__ float s_bias[2];
    AdamParams a = a_in;
    PeerParams pp = pp_in;
    const long long lo = (long long)blockIdx.x * a.chunk;
    const long long hi = min(a.n, lo + a.chunk);
    __shared__ unsigned int s_tag;
    __shared__ int s_peer_ok;
    long long dev_step = 0, dev_gen = 0;
    if (a.counters) {  // uniform branch: device-resident step / flag generation (graph capture)
        dev_step = *reinterpret_cast<volatile long long *>(a.counters) + 1;
        dev_gen = *reinterpret_cast<volatile long long *>(a.counters + 1) + 1;
        if (PEER) pp.gen = (unsigned int)dev_gen;
        if (threadIdx.x == 0) {  // torch's bias corrections, in double as Python forms them
            const double bc1 = 1.0 - pow(a.beta1, (double)dev_step), bc2 = 1.0 - pow(a.beta2_d, (double)dev_step);
            s_bias[0] = (float)(a.lr / bc1);
            s_bias[1] = (float)(1.0 / sqrt(bc2));
        }
    }
    if (PEER) {
        // "my gradients are final" to every rank (they were written by earlier kernels of this stream), then wait
        // for every rank's word in this rank's flag area
        if (threadIdx.x == 0) s_peer_ok = 1;
        __syncthreads();
        if (blockIdx.x == 0 && threadIdx.x < pp.world) st_release_sys_u32(pp.flags[threadIdx.x] + pp.rank, pp.gen);
        if (threadIdx.x < pp.world && !peer_wait(pp.flags[pp.rank] + threadIdx.x, pp.gen)) s_peer_ok = 0;
        __syncthreads();
        if (blockIdx.x == 0 && threadIdx.x < pp.n_extra) pp.extra_out[threadIdx.x] = load_grad1<true>(a, pp, a.n + threadIdx.x);
        if (pp.two_shot) {
            // reduce-scatter + all-gather of this CTA's chunk: this rank sums part `rank`, writes it to every rank
            const long long lo4 = lo / 4, cnt = hi / 4 - lo4;
            const long long ps = (cnt + pp.world - 1) / pp.world;
            const long long b = lo4 + pp.rank * ps, e = min(hi / 4, b + ps);
            if (pp.mc_g) {
                for (long long i = b + threadIdx.x; i < e; i += kOptThreads)
                    multimem_st_f4(reinterpret_cast<float4 *>(pp.mc_red) + i,
                                   multimem_ld_reduce_f4(reinterpret_cast<const float4 *>(pp.mc_g) + i));
            } else
            for (long long i = b + threadIdx.x; i < e; i += kOptThreads) {
                float4 v[kMaxPeers];
#pragma unroll
                for (int r = 0; r < kMaxPeers; ++r)
                    if (r < pp.world) v[r] = ld_sys_f4(reinterpret_cast<const float4 *>(pp.g[r]) + i);
                float4 sum = v[0];
#pragma unroll
                for (int r = 1; r < kMaxPeers; ++r)
                    if (r < pp.world) { sum.x += v[r].x; sum.y += v[r].y; sum.z += v[r].z; sum.w += v[r].w; }
#pragma unroll
                for (int r = 0; r < kMaxPeers; ++r)
                    if (r < pp.world) st_sys_f4(reinterpret_cast<float4 *>(pp.red[r]) + i, sum);
            }
            __threadfence_system();
            __syncthreads();
            if (threadIdx.x < pp.world)
                st_release_sys_u32(pp.cflags[threadIdx.x] + (long long)blockIdx.x * kMaxPeers + pp.rank, pp.gen);
            if (threadIdx.x < pp.world &&
                !peer_wait(pp.cflags[pp.rank] + (long long)blockIdx.x * kMaxPeers + threadIdx.x, pp.gen))
                s_peer_ok = 0;
            __syncthreads();
        }
    }
    if (threadIdx.x == 0) {
        unsigned int t = *reinterpret_cast<volatile unsigned int *>(a.gen) + 1u;
        s_tag = t ? t : 1u;  // 0 is the "never written" tag of a zeroed workspace
    }
    // ---- phase 1: partial sum of squares, fixed order ----
    float ss[1] = {0.f};
    float4 rp[kResident], rg[kResident], rm[kResident], rv[kResident];
    if (RESIDENT) {
        const float4 *p4 = reinterpret_cast<const float4 *>(a.p);
        const float4 *m4 = reinterpret_cast<const float4 *>(a.m), *v4 = reinterpret_cast<const float4 *>(a.v);
#pragma unroll
        for (int k = 0; k < kResident; ++k) {
            const long long i = lo / 4 + threadIdx.x + (long long)k * kOptThreads;
            if (i < hi / 4) { rg[k] = load_grad4<PEER>(a, pp, i); rp[k] = p4[i]; rm[k] = m4[i]; rv[k] = v4[i]; }
            else rg[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int k = 0; k < kResident; ++k)  // same per-thread order as the streaming path
            ss[0] += (rg[k].x * rg[k].x + rg[k].y * rg[k].y) + (rg[k].z * rg[k].z + rg[k].w * rg[k].w);
        for (long long i = (hi / 4) * 4 + threadIdx.x; i < hi; i += kOptThreads) {  // tail of n
            const float x = load_grad1<PEER>(a, pp, i);
            if (PEER) a.g[i] = x;
            ss[0] += x * x;
        }
    } else if (VEC) {
        float4 *g4 = reinterpret_cast<float4 *>(a.g);
        for (long long i = lo / 4 + threadIdx.x; i < hi / 4; i += kOptThreads) {
            const float4 x = load_grad4<PEER>(a, pp, i);
            if (PEER && !pp.two_shot) g4[i] = x;  // the reduced gradient, re-read by phase 2
            ss[0] += (x.x * x.x + x.y * x.y) + (x.z * x.z + x.w * x.w);
        }
        for (long long i = (hi / 4) * 4 + threadIdx.x; i < hi; i += kOptThreads) {  // tail of n
            const float x = load_grad1<PEER>(a, pp, i);
            if (PEER) a.g[i] = x;
            ss[0] += x * x;
        }
    } else {
        for (long long i = lo + threadIdx.x; i < hi; i += kOptThreads) {
            const float x = load_grad1<PEER>(a, pp, i);
            if (PEER) a.g[i] = x;
            ss[0] += x * x;
        }
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
    if (PEER && blockIdx.x == 0 && threadIdx.x < pp.world)  // every CTA of this rank has consumed its peer reads
        st_release_sys_u32(pp.flags[threadIdx.x] + kMaxPeers + pp.rank, pp.gen);
    const float norm = (PEER && !s_peer_ok) ? __int_as_float(0x7fc00000) : sqrtf(s_total);  // a peer never arrived: NaN
    float coef = 1.f;
    // torch clip_grad_norm_: clamp(max_norm / (norm + 1e-6), max=1).  torch.clamp PROPAGATES NaN, CUDA fminf returns the
    // non-NaN operand: a non-finite norm (NaN gradient, or the peer-timeout sentinel above) must poison the update
    // like the reference's does, not turn into coef = 1 with partial gradients
    if (a.max_norm > 0.f) coef = (norm == norm) ? fminf(a.max_norm / (norm + 1e-6f), 1.f) : norm;
    else if (PEER && !s_peer_ok) coef = norm;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *a.total_norm = norm;
        *reinterpret_cast<volatile unsigned int *>(a.gen) = tag;  // every CTA has read the old value (barrier above)
    }
    if (a.counters) {
        a.step_size = s_bias[0];      // written before the barriers above
        a.inv_bc2_sqrt = s_bias[1];
        if (blockIdx.x == 0 && threadIdx.x == 0) {  // every CTA has read the old values (grid barrier above)
            *reinterpret_cast<volatile long long *>(a.counters) = dev_step;
            if (PEER) *reinterpret_cast<volatile long long *>(a.counters + 1) = dev_gen;
        }
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
    if (PEER && blockIdx.x == 0 && threadIdx.x < pp.world)  // nobody still reads this rank's bucket once the kernel is done
        (void)peer_wait(pp.flags[pp.rank] + kMaxPeers + threadIdx.x, pp.gen);
}

// Shared host side of K6 (peer == nullptr) and K8.
static int32_t launch_clip_adam(const char *who, float *params, float *grads, float *exp_avg, float *exp_avg_sq, int64_t n,
                                float max_norm, double lr, double beta1, double beta2, double eps, int64_t step,
                                float *total_norm, void *workspace, void *stream, const PeerParams *peer,
                                int64_t *device_counters) {
    B200RL_REQUIRE(params && grads && exp_avg && exp_avg_sq && total_norm && workspace, -1, "%s: null pointer argument", who);
    B200RL_REQUIRE(n >= 1, -2, "%s: n=%lld", who, (long long)n);
    B200RL_REQUIRE(step >= 1 || device_counters, -2, "%s: step must be >= 1 (got %lld) unless device_counters is given", who, (long long)step);
    if (device_counters) step = 1;  // the host-side value is unused: the kernel forms the bias corrections itself
    B200RL_REQUIRE(lr >= 0 && beta1 >= 0 && beta1 < 1 && beta2 >= 0 && beta2 < 1 && eps >= 0, -2,
                   "%s: invalid hyper-parameters", who);
    static int max_grid_dev[kMaxDevices];      // co-resident CTAs of the streaming kernel, per device
    static int max_grid_res_dev[kMaxDevices];  // same for the register-resident instantiations
    int dev = 0;
    cudaGetDevice(&dev);
    const int di = (dev >= 0 && dev < kMaxDevices) ? dev : 0;
    int &max_grid = max_grid_dev[di];
    if (max_grid == 0) {
        int sms = 0, per_sm = 0, per_sm_res = 0, per_sm_peer = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, clip_adam_kernel<true, false, true>, kOptThreads, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_res, clip_adam_kernel<true, true, false>, kOptThreads, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_peer, clip_adam_kernel<true, true, true>, kOptThreads, 0);
        if (sms <= 0 || per_sm <= 0) {
            (void)cudaGetLastError();
            set_error("%s: occupancy query failed (no CUDA device?)", who);
            return (int32_t)cudaErrorUnknown;
        }
        max_grid = sms * (per_sm > 4 ? 4 : per_sm);
        if (max_grid > kMaxOptCtas) max_grid = kMaxOptCtas;
        if (per_sm_peer < per_sm_res) per_sm_res = per_sm_peer;
        max_grid_res_dev[di] = sms * (per_sm_res > 0 ? per_sm_res : 0);
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
    a.counters = reinterpret_cast<long long *>(device_counters);
    a.lr = lr; a.beta1 = beta1; a.beta2_d = beta2;
    a.total_norm = total_norm;
    a.gen = reinterpret_cast<unsigned int *>(reinterpret_cast<char *>(workspace) + kOptOffset);
    a.words = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(workspace) + kOptOffset + 64);
    uintptr_t align = reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads) |
                      reinterpret_cast<uintptr_t>(exp_avg) | reinterpret_cast<uintptr_t>(exp_avg_sq);
    PeerParams pp = {};
    if (peer) {
        pp = *peer;
        for (int r = 0; r < pp.world; ++r) align |= reinterpret_cast<uintptr_t>(pp.g[r]);
    }
    const bool vec = (align % 16) == 0;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kOptThreads);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;  // co-residency: the in-kernel grid barrier cannot deadlock
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const bool resident = vec && chunk <= (long long)kResident * 4 * kOptThreads && grid <= max_grid_res_dev[di] &&
                          env_int("B200RL_K6_RESIDENT", 1);
    cudaError_t e;
    if (peer)
        e = resident ? cudaLaunchKernelEx(&cfg, clip_adam_kernel<true, true, true>, a, pp)
            : vec    ? cudaLaunchKernelEx(&cfg, clip_adam_kernel<true, false, true>, a, pp)
                     : cudaLaunchKernelEx(&cfg, clip_adam_kernel<false, false, true>, a, pp);
    else
        e = resident ? cudaLaunchKernelEx(&cfg, clip_adam_kernel<true, true, false>, a, pp)
            : vec    ? cudaLaunchKernelEx(&cfg, clip_adam_kernel<true, false, false>, a, pp)
                     : cudaLaunchKernelEx(&cfg, clip_adam_kernel<false, false, false>, a, pp);
    if (e != cudaSuccess) {
        set_error("clip_adam_kernel launch: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    return check_launch("clip_adam_kernel");
}

}  // namespace b200rl

extern "C" int32_t b200rl_clip_adam_step(float *params, float *grads, float *exp_avg, float *exp_avg_sq,
                                         int64_t n, float max_norm, double lr, double beta1, double beta2,
                                         double eps, int64_t step, int64_t *device_counters, float *total_norm,
                                         void *workspace, void *stream) {
    return b200rl::launch_clip_adam("clip_adam_step", params, grads, exp_avg, exp_avg_sq, n, max_norm, lr, beta1, beta2, eps,
                                    step, total_norm, workspace, stream, nullptr, device_counters);
}

extern "C" int32_t b200rl_peer_allreduce_clip_adam_step(
    float *params, const float *const *peer_buckets, uint32_t *const *peer_flags, float *const *peer_reduced,
    uint32_t *const *peer_cta_flags, const float *mc_buckets, float *mc_reduced, int32_t world, int32_t rank,
    float *grads_out, float *extra_out, int64_t n_extra, float *exp_avg, float *exp_avg_sq, int64_t n, float max_norm,
    double lr, double beta1, double beta2, double eps, int64_t step, int64_t flag_generation,
    int64_t *device_counters, float *total_norm, void *workspace, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(peer_buckets && peer_flags && grads_out, -1, "peer_allreduce_clip_adam_step: null pointer argument");
    B200RL_REQUIRE(device_counters || (flag_generation >= 1 && flag_generation < (1ll << 31)), -2,
                   "peer_allreduce_clip_adam_step: flag_generation=%lld (a counter >= 1 that every rank advances by one per call)",
                   (long long)flag_generation);
    B200RL_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, -2,
                   "peer_allreduce_clip_adam_step: world=%d rank=%d (at most %d ranks)", world, rank, kMaxPeers);
    B200RL_REQUIRE(n_extra >= 0 && n_extra <= kOptThreads && (n_extra == 0 || extra_out), -2,
                   "peer_allreduce_clip_adam_step: n_extra=%lld", (long long)n_extra);
    B200RL_REQUIRE(device_counters || (step >= 1 && step < (1ll << 31)), -2, "peer_allreduce_clip_adam_step: step=%lld", (long long)step);
    PeerParams pp = {};
    for (int r = 0; r < world; ++r) {
        B200RL_REQUIRE(peer_buckets[r] && peer_flags[r], -1, "peer_allreduce_clip_adam_step: rank %d pointer is null", r);
        B200RL_REQUIRE(peer_buckets[r] != grads_out, -2, "peer_allreduce_clip_adam_step: grads_out must not alias a bucket");
        pp.g[r] = peer_buckets[r];
        pp.flags[r] = peer_flags[r];
    }
    // route: B200RL_K8_TWO_SHOT=1/0 forces it, default two-shot for more than 2 ranks (needs the shared reduced buffers)
    const int want = env_int("B200RL_K8_TWO_SHOT", -1);
    pp.two_shot = (peer_reduced && peer_cta_flags && (want == 1 || (want < 0 && world > 2))) ? 1 : 0;
    if (pp.two_shot) {
        uintptr_t al = 0;
        for (int r = 0; r < world; ++r) {
            B200RL_REQUIRE(peer_reduced[r] && peer_cta_flags[r], -1, "peer_allreduce_clip_adam_step: rank %d pointer is null", r);
            pp.red[r] = peer_reduced[r];
            pp.cflags[r] = peer_cta_flags[r];
            al |= reinterpret_cast<uintptr_t>(peer_reduced[r]) | reinterpret_cast<uintptr_t>(peer_buckets[r]);
        }
        B200RL_REQUIRE(peer_reduced[rank] == grads_out, -2,
                       "peer_allreduce_clip_adam_step: grads_out must be this rank's shared reduced buffer");
        if (al % 16 || ((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(exp_avg) |
                         reinterpret_cast<uintptr_t>(exp_avg_sq)) % 16))
            pp.two_shot = 0;  // unaligned buffers take the scalar one-shot route
    }
    if (pp.two_shot && mc_buckets && mc_reduced && env_int("B200RL_K8_NVLS", 1) &&
        ((reinterpret_cast<uintptr_t>(mc_buckets) | reinterpret_cast<uintptr_t>(mc_reduced)) % 16) == 0) {
        pp.mc_g = mc_buckets;
        pp.mc_red = mc_reduced;
    }
    pp.extra_out = extra_out;
    pp.world = world; pp.rank = rank; pp.n_extra = (int)n_extra;
    pp.gen = (unsigned int)flag_generation;  // owned by the flag areas (the bucket), NOT the optimizer's step: see b200rl.h
    return launch_clip_adam("peer_allreduce_clip_adam_step", params, grads_out, exp_avg, exp_avg_sq, n, max_norm, lr, beta1,
                            beta2, eps, step, total_norm, workspace, stream, &pp, device_counters);
}
