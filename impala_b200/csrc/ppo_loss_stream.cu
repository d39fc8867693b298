// K3, streaming variant: fused PPO clipped-surrogate loss forward + backward as a PERSISTENT, ring-pipelined kernel.
// Replaces reference losses.py:134-155 (after model(s)) and the loss part of loss.backward() at
// agents/ppo/learning.py:134.  Same arithmetic as the tile kernel (ppo_loss.cu), another schedule.
//
// The tile kernel gives every CTA one 256-sample tile: load, compute, store, reduce -- 8192 CTAs at BASELINE config 3, each
// a serial chain, 18 waves: 0.64 of the HBM peak when a learner step issues it alone.  Here the grid is two CTAs per SM and
// CTA k streams its contiguous share of the minibatch through a ring of S shared-memory stages:
//   * a PRODUCER warp keeps S chunks (256 samples: logits + behaviour logits, two TMA bulk copies, SASS UBLKCP) in flight,
//     sends a finished chunk's gradients out with one bulk store and refills the stage when the store has read it;
//   * 8 CONSUMER warps each own 32 samples of every chunk and run at their own pace -- full[stage] (TMA complete_tx) and
//     done[stage] (one arrive per consumer warp) mbarriers are the only synchronisation, there is no block-wide barrier;
//   * the six loss sums stay in registers over the CTA's whole share; one reduction per CTA at the end (last-arriving warp),
//     published as generation-tagged words to one collector warp (warp_collect, common.cuh).
// Measured (T=256, B=8192, A=15, K2 + K3 per step): 108.0 us with the tile kernel, 88.9 us with one CTA per SM and 4 stages,
// 75.6 us with two CTAs per SM and 2 stages each (the default; three CTAs per SM: 82 us).
// 200 B/sample at A=15: every input byte read once, every gradient byte written once.
#include "ppo_loss.cuh"
#include "softmax.cuh"

namespace b200rl {

constexpr int kStreamChunk = 256;      // samples per chunk = 8 consumer warps x 32 lanes
constexpr int kStreamConsumers = 8;
constexpr int kStreamThreads = (kStreamConsumers + 1) * 32;

struct PpoStreamParams {
    PpoParams p;
    int S;                       // stages
    int chunks_lo, chunks_rem;   // CTA b owns chunks_lo + (b < chunks_rem) chunks starting at b * chunks_lo + min(b, chunks_rem)
    int stage_words;             // 2 * kStreamChunk * A
};

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int AT, bool EXACT>
__global__ void __launch_bounds__(kStreamThreads, 2) ppo_loss_stream_kernel(const __grid_constant__ PpoStreamParams q) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const PpoParams &p = q.p;
    const int A = EXACT ? AT : p.A, S = q.S;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *smem = reinterpret_cast<float *>(smem_raw);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)S * q.stage_words);
    uint64_t *done = full + S;
    float *red = reinterpret_cast<float *>(done + S);  // [6][kStreamConsumers]
    int *ctl = reinterpret_cast<int *>(red + 6 * kStreamConsumers);

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(full + s, 1); mbar_init(done + s, kStreamConsumers); }
        ctl[0] = 0;
        fence_mbar_init();
    }
    pdl_launch_dependents();
    const bool late_wait = p.flags & B200RL_INPUTS_STABLE;
    unsigned int gen = 0;
    const int b = (int)blockIdx.x;
    const long long c_begin = (long long)b * q.chunks_lo + min(b, q.chunks_rem);
    const int nc = q.chunks_lo + (b < q.chunks_rem ? 1 : 0);
    auto chunk_n0 = [&](int c) { return (c_begin + c) * (long long)kStreamChunk; };
    auto chunk_count = [&](int c) { return (int)min((long long)kStreamChunk, p.N - chunk_n0(c)); };
    if (!late_wait) {
        // A HINT, not a read: while the preceding kernel (normally the lambda-return scan that writes v_target) still runs,
        // ask L2 for the chunks this CTA loads first (cp.async.bulk.prefetch.L2); every value is read after the wait below.
        if (warp == kStreamConsumers && lane < min(S, nc)) {
            const size_t g0 = (size_t)chunk_n0(lane) * A;
            const uint32_t bytes = (uint32_t)(chunk_count(lane) * A * 4);
            bulk_prefetch_l2(p.logits + g0, bytes);
            bulk_prefetch_l2(p.pi_ref + g0, bytes);
        }
        pdl_wait();  // inputs may come from the preceding kernel
        gen = grid_generation(p.workspace);
    }
    __syncthreads();

    if (warp == kStreamConsumers) {
        // ---- producer: one thread drives the ring ------------------------------------------------
        if (lane == 0) {
            auto issue_load = [&](int c) {
                const int st = c % S;
                float *sx = smem + (size_t)st * q.stage_words;
                const size_t g0 = (size_t)chunk_n0(c) * A;
                const uint32_t bytes = (uint32_t)(chunk_count(c) * A * 4);
                mbar_expect_tx(full + st, 2u * bytes);
                bulk_g2s(sx, p.logits + g0, bytes, full + st);
                bulk_g2s(sx + (size_t)kStreamChunk * A, p.pi_ref + g0, bytes, full + st);
            };
            for (int c = 0; c < min(S, nc); ++c) issue_load(c);
            for (int c = 0; c < nc; ++c) {
                const int st = c % S;
                mbar_wait(done + st, (uint32_t)(c / S) & 1u);  // every consumer warp has written its gradient rows
                bulk_s2g(p.dlogits + (size_t)chunk_n0(c) * A, smem + (size_t)st * q.stage_words, (uint32_t)(chunk_count(c) * A * 4));
                bulk_commit();
                if (c + S < nc) {  // refill the stage once the store has read it
                    bulk_wait_read0();
                    issue_load(c + S);
                }
            }
            bulk_wait_read0();  // shared memory must outlive the bulk stores' reads
        }
        return;
    }

    // ---- consumers: warp w owns samples [32 w, 32 w + 32) of every chunk ------------------------------
    const int s = warp * 32 + lane;
    float sums[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // pg, td, entropy, kl, ratio, target
    long long act = 0;
    float vt = 0.f, v = 0.f;
    auto fetch_small = [&](int c) {  // the 16 bytes per sample outside the logits: plain loads, one chunk ahead
        if (c >= nc || s >= chunk_count(c)) return;
        const long long n = chunk_n0(c) + s;
        act = __ldg(p.actions + n);
        vt = __ldg(p.v_target + n);
        v = __ldg(p.v + n);
    };
    fetch_small(0);
    const float lo = 1.f - p.clip_coeff, hi = 1.f + p.clip_coeff;
    for (int c = 0; c < nc; ++c) {
        const int st = c % S;
        float *sx = smem + (size_t)st * q.stage_words, *sy = sx + (size_t)kStreamChunk * A;
        const bool valid = s < chunk_count(c);
        const long long n = chunk_n0(c) + s;
        mbar_wait(full + st, (uint32_t)(c / S) & 1u);
        const long long c_act = act;
        const float c_vt = vt, c_v = v;
        if (valid) {
            SampleStats<AT> stt;
            sample_stats<AT, EXACT>(stt, sx + s * A, sy + s * A, A, c_act);
            fetch_small(c + 1);
            const float ratio = stt.rho;
            const float adv = c_vt - c_v;                               // losses.py:137
            const float pg1 = -(adv * ratio);                           // :140
            const float pg2 = -fminf(fmaxf(ratio, lo), hi) * adv;       // :141
            sums[0] += fmaxf(pg1, pg2);                                 // :142
            sums[1] += adv * adv;
            sums[2] += stt.H;
            sums[3] += stt.kl;
            sums[4] += ratio;
            sums[5] += c_vt;
            // torch.max splits ties evenly and clamp passes gradient on [lo,hi]: inside the
            // range both branches carry d/d ratio = -adv (halves add up); outside only pg1 can.
            const bool inside = ratio >= lo && ratio <= hi;
            const float g_ratio = (inside || pg1 > pg2) ? -adv : 0.f;
            const float g_lp = g_ratio * ratio * p.inv_n;               // d ratio / d lp = ratio
            p.dv[n] = -adv * p.inv_n;                                   // d(0.5 adv^2)/dv, adv detached in pg
            write_dlogits<AT, EXACT>(stt, sx + s * A, A, g_lp, p.entropy_cost * p.inv_n);
        } else {
            fetch_small(c + 1);
        }
        fence_proxy_async_smem();  // gradient rows -> visible to the producer's bulk store
        __syncwarp();
        if (lane == 0) mbar_arrive(done + st);
    }

    // ---- scalars: the CTA's last-arriving consumer warp adds the warps' sums in index order and publishes ------
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int k = 0; k < 6; ++k) sums[k] += __shfl_xor_sync(0xffffffffu, sums[k], o);
    }
    int arrived = 0;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 6; ++k) red[k * kStreamConsumers + warp] = sums[k];
        __threadfence_block();
        arrived = atomicAdd(&ctl[0], 1);
    }
    arrived = __shfl_sync(0xffffffffu, arrived, 0);
    if (arrived != kStreamConsumers - 1) return;
    __threadfence_block();
    float tot[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) tot[k] = lane < kStreamConsumers ? red[k * kStreamConsumers + lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int k = 0; k < 6; ++k) tot[k] += __shfl_xor_sync(0xffffffffu, tot[k], o);
    }
    if (late_wait) {
        pdl_wait();  // the workspace and scalars are shared with the previous launch
        gen = grid_generation(p.workspace);
    }
    if (blockIdx.x != gridDim.x - 1) {
        warp_publish<6>(tot, p.workspace, gen);
        return;
    }
    const float inv_n = p.inv_n, ec = p.entropy_cost;
    float *scalars = p.scalars;
    warp_collect<6>(tot, p.workspace, gen, [=](const float(&t)[6]) {
        const float pg = t[0] * inv_n, td = 0.5f * t[1] * inv_n, ent = t[2] * inv_n;
        scalars[B200RL_S_LOSS] = pg + td - ec * ent;  // losses.py:146
        scalars[B200RL_S_ENTROPY] = ent;
        scalars[B200RL_S_TD] = td;
        scalars[B200RL_S_PG] = pg;
        scalars[B200RL_S_KL] = t[3] * inv_n;
        scalars[B200RL_S_RATIO] = t[4] * inv_n;
        scalars[B200RL_S_TARGET] = t[5] * inv_n;
        scalars[7] = 0.f;
    });
}

template <int AT, bool EXACT>
static int32_t launch_ppo_stream_t(const PpoStreamParams &q, int grid, size_t smem, cudaStream_t stream) {
    static size_t smem_cache[kMaxDevices];
    if (int32_t rc = ensure_dynamic_smem(ppo_loss_stream_kernel<AT, EXACT>, smem, smem_cache, "ppo_loss")) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kStreamThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // PDL: see pdl_wait() in the kernel
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = env_int("B200RL_PDL", 1) ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&cfg, ppo_loss_stream_kernel<AT, EXACT>, q);
    if (e != cudaSuccess) {
        set_error("ppo_loss_stream_kernel launch: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    return check_launch("ppo_loss_stream_kernel");
}

// returns 1 when the shape is not taken (the caller then uses the tile kernel)
int32_t launch_ppo_stream(const PpoParams &p, cudaStream_t stream) {
    // TMA bulk copies: 16-byte aligned bases; chunk starts are multiples of 256 samples, the tail needs N * A % 4 == 0
    if (((reinterpret_cast<uintptr_t>(p.logits) | reinterpret_cast<uintptr_t>(p.pi_ref) |
          reinterpret_cast<uintptr_t>(p.dlogits)) % 16) != 0 || (p.N * p.A) % 4 != 0)
        return 1;
    const long long n_chunks = (p.N + kStreamChunk - 1) / kStreamChunk;
    int dev = 0;
    cudaGetDevice(&dev);
    static int sm_cache[kMaxDevices];
    const int di = (dev >= 0 && dev < kMaxDevices) ? dev : 0;
    if (!sm_cache[di]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        sm_cache[di] = n;
    }
    const int sms = sm_cache[di];
    const int cps = max(1, min(3, env_int("B200RL_K3_CTAS_PER_SM", 2)));   // CTAs per SM the grid is sized for
    const int grid = sms * cps;
    if (n_chunks < 2LL * grid) return 1;  // small minibatches: the tile kernel's many CTAs fill the machine better
    if (grid > kMaxPartialCtas) return 1;
    PpoStreamParams q;
    q.p = p;
    q.chunks_lo = (int)(n_chunks / grid);
    q.chunks_rem = (int)(n_chunks % grid);
    q.stage_words = 2 * kStreamChunk * p.A;
    const size_t stage_bytes = (size_t)q.stage_words * 4;
    int S = env_int("B200RL_K3_STAGES", 2);
    S = max(2, min(S, (int)((200 * 1024 / cps) / stage_bytes)));
    S = min(S, q.chunks_lo + 1);
    q.S = S;
    const size_t smem = stage_bytes * S + (size_t)S * 16 + 6 * kStreamConsumers * 4 + 16 + 128;
    switch (p.A) {
        case 4: return launch_ppo_stream_t<4, true>(q, grid, smem, stream);
        case 6: return launch_ppo_stream_t<6, true>(q, grid, smem, stream);
        case 15: return launch_ppo_stream_t<15, true>(q, grid, smem, stream);
        case 18: return launch_ppo_stream_t<18, true>(q, grid, smem, stream);
        default: break;
    }
    if (p.A <= 8) return launch_ppo_stream_t<8, false>(q, grid, smem, stream);
    if (p.A <= 16) return launch_ppo_stream_t<16, false>(q, grid, smem, stream);
    return launch_ppo_stream_t<32, false>(q, grid, smem, stream);
}

}  // namespace b200rl
