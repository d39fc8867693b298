// K1: fused V-trace + IMPALA loss, forward and backward in ONE launch -- persistent, warp-group pipelined.
// Replaces reference agents/impala/learning.py:146-160,167-168 (+ rlego V-trace, learning.py:15-26).
//
// One learner step issues ONE K1, so a single launch has to keep HBM busy by itself.  The first version
// (impala_loss_tile.cu) gave every CTA one tile: all CTAs loaded, then all computed, then all stored, and
// every CTA paid a block-wide reduction.  Here the grid is one CTA per SM; CTA k owns a CONTIGUOUS range of
// batch columns, cut into UNITS of C whole columns (C*T <= 32*W samples, one thread per sample; T = 20:
// W = 2 warps, C = 3) and into CHUNKS of GPC consecutive units.
//
//   * loads: per chunk TWO TMA bulk copies (cp.async.bulk, SASS UBLKCP) -- logits and behaviour logits --
//     on the chunk's mbarrier, all chunks issued the moment the kernel starts.  Measured on B200
//     (scripts/micro/tma_lat.cu, profiles/r02_tma_bulk_latency.txt): a bulk copy costs the SM's TMA unit
//     ~150 cycles whatever its size and copies complete in issue order, so FEW LARGE copies per SM land
//     one after the other over the load time; many small ones (a copy per tensor per unit, the first
//     attempt: 55 per SM) are op-rate bound at a third of the bandwidth.  The 20 bytes per sample of
//     actions / rewards / discounts / values are plain coalesced loads of the owning thread;
//   * a GROUP of W warps takes the units g, g + G, ... of its CTA and runs as soon as the unit's chunk has
//     landed, synchronising only with itself (named barrier / __syncwarp): softmax statistics per sample
//     (registers), the (T-1)-step V-trace recursion as a register chain in C threads (reference operation
//     order), closed-form gradients written over the staged logits.  Groups are in different phases at
//     any time, so loads, arithmetic and stores of different chunks overlap inside the launch;
//   * stores: the group that finishes a chunk last sends it out with ONE bulk store and, when the CTA has
//     more chunks than stages, refills the stage;
//   * the five loss sums stay in registers across a thread's units; the CTA's last-arriving warp adds the
//     warps' sums (no block-wide barrier) and publishes them as generation-tagged words; one collector
//     warp sums the CTAs in a fixed order (warp_collect, common.cuh).
//
// Every input byte is read once, every gradient byte written once: 204 B/sample at A=15.
#include "impala_loss.cuh"
#include "softmax.cuh"

namespace b200rl {

struct K1Params {
    const float *logits, *pi_ref, *rewards, *discounts, *values;
    const long long *actions;
    float *dlogits, *dvalues, *scalars;
    void *workspace;
    int B, T, A;
    int C, W, G, U;             // columns per unit, warps per group, groups per CTA, units
    int GPC, NSTAGE;            // units per chunk, chunk stages in shared memory
    int unit_words, ns4;        // C*T*A, C*T rounded up to 4 (4-byte words)
    unsigned int inv_T;         // ceil(2^32 / T)
    // no integer division on the device (a 64-bit one is ~100 instructions in every warp): the units-per-CTA split and
    // the magic reciprocals ceil(2^32 / d) of W, GPC, NSTAGE come from the host (exact for operands < 2^16)
    int units_lo, units_rem;    // CTA b owns units_lo + (b < units_rem) units starting at b * units_lo + min(b, units_rem)
    unsigned int inv_W, inv_GPC, inv_NSTAGE;
    float lambda_, one_minus_lambda, clip_rho, clip_pg_rho, entropy_coeff, inv_n_pg, inv_n_all;
    uint32_t flags;
    int prefetch;               // L2 prefetch hint of the CTA's inputs before the PDL dependency wait
    int debug;                  // B200RL_K1_DEBUG (profiling): 1 = skip the arithmetic, 2 = skip the grid collector, 4 = skip the small-field loads, 8 = reduce after the gradients
};

// x / d for small non-negative x with the host's magic m = ceil(2^32 / d) (d = 1 has no 32-bit magic)
__device__ __forceinline__ int fast_div(int x, int d, unsigned int m) {
    return d == 1 ? x : (int)__umulhi((unsigned)x, m);
}

__device__ __forceinline__ void group_sync(int id, int nthreads) {
    if (nthreads == 32) __syncwarp();
    else asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// AT <= 16: up to 20 warps per CTA at <= 96 registers; wider action spaces: 10 warps
template <int AT> constexpr int k1_max_threads() { return AT <= 16 ? 640 : 320; }

// TT, CW, CC, CG: compile-time rollout length T, warps per group W, columns per unit C and units per chunk GPC (0 = the
// run-time values of the parameter block).  The kernel is issue-bound once its data is in flight (ncu: ~940 warp
// instructions per warp in the all-run-time version, ~260 of them softmax arithmetic): with the reference's rollout
// length as a constant the index arithmetic folds away and the recursion unrolls, as the exact-A instantiations do for
// the logit loops.  Other shapes take the run-time instantiation.
template <int AT, bool EXACT, bool TRACE, int TT, int CW, int CC, int CG>
__global__ void __launch_bounds__(k1_max_threads<AT>(), 1) impala_loss_kernel(const __grid_constant__ K1Params p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int T = TT ? TT : p.T, A = EXACT ? AT : p.A, C = CC ? CC : p.C, G = p.G, GPC = CG ? CG : p.GPC, NSTAGE = p.NSTAGE;
    const int W = CW ? CW : p.W;
    const int unit_words = (TT && EXACT) ? CC * TT * AT : p.unit_words, ns4 = TT ? ((CC * TT + 3) & ~3) : p.ns4;
    const int gthreads = W * 32;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    const int grp = CW ? warp / CW : fast_div(warp, p.W, p.inv_W);
    const int i = threadIdx.x - grp * gthreads;  // thread (= sample) index inside the group
    const int bar_id = grp + 1;                  // named barrier of the group (0 is __syncthreads)
    const int chunk_words = GPC * unit_words;
    // shared memory: [NSTAGE][logits chunk | pi_ref chunk] | per-group scratch (td, g*c), e, nx | mbarriers | done | red | ctl
    float *smem = reinterpret_cast<float *>(smem_raw);
    float *scratch = smem + (size_t)NSTAGE * 2 * chunk_words + (size_t)grp * 4 * ns4;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)NSTAGE * 2 * chunk_words + (size_t)G * 4 * ns4);
    int *done = reinterpret_cast<int *>(bars + NSTAGE);  // [NSTAGE] groups that finished the stage's chunk
    int *issued = done + NSTAGE;                         // [NSTAGE] highest round whose load has been issued into the stage
    float *red = reinterpret_cast<float *>(issued + NSTAGE);  // [5][nwarps]
    int *ctl = reinterpret_cast<int *>(red + 5 * nwarps);   // [0] warps arrived at the reduction

    const bool trace = TRACE && i == 0 && (int)(blockIdx.x * G + grp) < kMaxTraceCtas;
    unsigned long long *tr = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(p.workspace) + kTraceOffset) +
                             (size_t)(blockIdx.x * G + grp) * 16;
#define K1_STAMP(k) do { if (TRACE && trace) { tr[k] = globaltimer_ns(); tr[8 + (k)] = (unsigned long long)clock64(); } } while (0)
    if (trace) { unsigned int smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid)); tr[7] = smid; }

    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(bars + s, 1); done[s] = 0; issued[s] = 0; }
        ctl[0] = 0;
        fence_mbar_init();
    }
    const unsigned long long t_entry = (TRACE && threadIdx.x == 0) ? globaltimer_ns() : 0ull;
    pdl_launch_dependents();
    const bool late_wait = p.flags & B200RL_INPUTS_STABLE;
    unsigned int gen = 0;
    if (!late_wait) {
        if (p.prefetch) {
            // A HINT, not a read: while the preceding kernel is still running, ask L2 for this CTA's share of the
            // inputs (cp.async.bulk.prefetch.L2).  Address translation and the HBM reads then overlap the
            // predecessor's tail; L2 is the coherence point, so lines the predecessor still writes are simply
            // updated there, and every value is read only after the dependency wait below.
            if (lane == 0 && warp < 4) {
                const int b = (int)blockIdx.x, us = b * p.units_lo + min(b, p.units_rem), ue = us + p.units_lo + (b < p.units_rem ? 1 : 0);
                const size_t e0 = (size_t)us * C * T, e1 = (size_t)min(ue * C, p.B) * T;   // samples [e0, e1)
                const uint32_t ns = (uint32_t)(e1 - e0), nb = ns * (uint32_t)A * 4u;
                const int w = nwarps >= 4 ? warp : -1;   // four issuing warps when the CTA has them, else warp 0 does all
                if (ns && (w == 0 || (w < 0 && warp == 0))) bulk_prefetch_l2(p.logits + e0 * A, nb);
                if (ns && (w == 1 || (w < 0 && warp == 0))) bulk_prefetch_l2(p.pi_ref + e0 * A, nb);
                if (ns && (w == 2 || (w < 0 && warp == 0))) {
                    bulk_prefetch_l2(p.actions + e0, ns * 8u);
                    bulk_prefetch_l2(p.rewards + e0, ns * 4u);
                }
                if (ns && (w == 3 || (w < 0 && warp == 0))) {
                    bulk_prefetch_l2(p.discounts + e0, ns * 4u);
                    bulk_prefetch_l2(p.values + e0, ns * 4u);
                }
            }
        }
        pdl_wait();  // inputs may come from the preceding kernel
        gen = grid_generation(p.workspace);  // consumed by the last warp at the very end: latency hidden
    }
    // CTA-level stamps of the trace instantiation: rows 2048 + 1024 * (generation & 1) + blockIdx (two consecutive launches
    // stay apart): 0 last warp at the reduction, 1 published, 2 kernel entry, 3 past the dependency wait, 4 last instruction
    unsigned long long *ctr = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(p.workspace) + kTraceOffset) +
                              (size_t)(2048 + 1024 * (gen & 1u) + (blockIdx.x & 1023u)) * 16;
    if (TRACE && threadIdx.x == 0) { ctr[2] = t_entry; ctr[3] = globaltimer_ns(); }
    __syncthreads();
    K1_STAMP(0);

    // this CTA's contiguous range of units, its chunks
    const int ustart = (int)blockIdx.x * p.units_lo + min((int)blockIdx.x, p.units_rem);
    const int n_units = p.units_lo + ((int)blockIdx.x < p.units_rem ? 1 : 0);
    const int n_chunks = CG ? (n_units + CG - 1) / CG : fast_div(n_units + GPC - 1, GPC, p.inv_GPC);
    auto stage_of = [&](int c, int &round) { round = fast_div(c, NSTAGE, p.inv_NSTAGE); return c - round * NSTAGE; };
    auto chunk_units = [&](int c) { return min(GPC, n_units - c * GPC); };
    auto chunk_col0 = [&](int c) { return (ustart + c * GPC) * C; };
    auto chunk_cols = [&](int c) { return min((ustart + c * GPC + chunk_units(c)) * C, p.B) - chunk_col0(c); };
    auto issue_chunk = [&](int c) {  // one thread
        int round;
        const int st = stage_of(c, round);
        float *sx = smem + (size_t)st * 2 * chunk_words;
        const size_t e0 = (size_t)chunk_col0(c) * T * A;
        const uint32_t bytes = (uint32_t)(chunk_cols(c) * T * A * 4);
        mbar_expect_tx(bars + st, 2u * bytes);
        bulk_g2s(sx, p.logits + e0, bytes, bars + st);
        bulk_g2s(sx + chunk_words, p.pi_ref + e0, bytes, bars + st);
    };
    if (lane == 0)  // copies complete in issue order: chunk 0 first.  Spread over warps: an issue costs ~55 cycles
        for (int c = warp; c < min(NSTAGE, n_chunks); c += nwarps) issue_chunk(c);

    const bool detach_adv = p.flags & B200RL_DETACH_ADV, detach_rho = p.flags & B200RL_DETACH_RHO;
    const int col = TT ? i / TT : (int)__umulhi((unsigned)i, p.inv_T), t = i - col * T;
    float2 *s_tg = reinterpret_cast<float2 *>(scratch);
    float *s_e = scratch + 2 * ns4, *s_nx = s_e + ns4;
    float sums[5] = {0.f, 0.f, 0.f, 0.f, 0.f};  // pg, value, entropy, kl, ratio

    // the 20 bytes per sample outside the logits: plain loads, fetched one unit ahead
    long long act = 0;
    float rew = 0.f, disc = 0.f, v = 0.f, v_next = 0.f;
    auto fetch_small = [&](int j) {
        if (j >= n_units || (p.debug & 4)) return;
        const int b0 = (ustart + j) * C, nsu = min(C, p.B - b0) * T;
        const size_t n = (size_t)b0 * T + i;
        if (i < nsu) {
            act = __ldg(p.actions + n);
            rew = __ldg(p.rewards + n);
            disc = __ldg(p.discounts + n);
            v = __ldg(p.values + n);
            if (t < T - 1) v_next = __ldg(p.values + n + 1);
        }
    };
    fetch_small(grp);

    // One reduction per CTA by its last-arriving warp, no block-wide barrier: a warp leaves its five sums in shared
    // memory; the warp that arrives last adds the warps' sums (fixed order everywhere: bit-reproducible scalars) and
    // publishes the CTA's partials (warp_publish).  In the collector CTA that warp only notes its duty: it polls for the
    // other CTAs' words at the very end, when they have long landed.
    bool reduced = false, is_collector = false;
    auto cta_reduce_publish = [&]() {  // consumes `sums`
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int k = 0; k < 5; ++k) sums[k] += __shfl_xor_sync(0xffffffffu, sums[k], o);
        }
        int arrived = 0;
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < 5; ++k) red[k * nwarps + warp] = sums[k];
            __threadfence_block();
            arrived = atomicAdd(&ctl[0], 1);
        }
        arrived = __shfl_sync(0xffffffffu, arrived, 0);
        if (arrived != nwarps - 1) return;
        if (TRACE && lane == 0) ctr[0] = globaltimer_ns();
        __threadfence_block();
        // lane w takes warp w's sums; xor-shuffle tree in a fixed order (nwarps <= 32)
        float cta_tot[5];
#pragma unroll
        for (int k = 0; k < 5; ++k) cta_tot[k] = lane < nwarps ? red[k * nwarps + lane] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int k = 0; k < 5; ++k) cta_tot[k] += __shfl_xor_sync(0xffffffffu, cta_tot[k], o);
        }
        if (late_wait) {
            pdl_wait();  // the workspace and scalars are shared with the previous launch
            gen = grid_generation(p.workspace);
        }
        if (blockIdx.x == gridDim.x - 1) {  // the collector parks its own totals in shared memory until the end
            is_collector = true;
            __syncwarp();
#pragma unroll
            for (int k = 0; k < 5; ++k)
                if (lane == k) red[k] = cta_tot[k];
        } else if (!(p.debug & 2)) warp_publish<5>(cta_tot, p.workspace, gen);
        if (TRACE && lane == 0) ctr[1] = globaltimer_ns();
    };

    bool first = true;
    for (int j = grp; j < n_units; j += G) {
        const int c = CG ? j / CG : fast_div(j, GPC, p.inv_GPC);
        int round;
        const int st_i = stage_of(c, round);
        float *sx = smem + (size_t)st_i * 2 * chunk_words + (size_t)(j - c * GPC) * unit_words;
        float *sy = sx + chunk_words;
        const int b0 = (ustart + j) * C, ucols = min(C, p.B - b0), nsu = ucols * T;
        const size_t n0 = (size_t)b0 * T;

        // Ring rounds: a chunk is completed -- and its stage refilled -- by the GPC groups that own its units only, so a group
        // can be TWO rounds ahead of the groups that still work on the stage's round-0 chunk.  A parity wait alone would then
        // match the long-completed round 0 instead of round 2.  So: first wait until the load of THIS round has been issued
        // (a shared-memory counter written after the issue), then the phase wait is unambiguous -- round r+1 cannot be issued
        // before this group has finished its unit of round r.
        if (round > 0)
            while (*reinterpret_cast<volatile int *>(issued + st_i) < round) {}
        mbar_wait(bars + st_i, (uint32_t)round & 1u);  // the unit's chunk has landed
        if (first) K1_STAMP(1);

        // ---- per-sample softmax statistics, TD error, trace coefficient ------------------------
        const bool valid = i < nsu && !(p.debug & 1);
        const bool has_next = valid && t < T - 1;
        SampleStats<AT> st;
        float c_pg = 0.f;
        const float u_rew = rew, u_disc = disc, u_v = v, u_vn = v_next;
        if (valid) {
            sample_stats<AT, EXACT>(st, sx + i * A, sy + i * A, A, act);
            c_pg = fminf(st.rho, p.clip_pg_rho);
            if (has_next) {
                s_tg[i] = make_float2(fminf(st.rho, p.clip_rho) * (u_rew + u_disc * u_vn - u_v),
                                      u_disc * (fminf(st.rho, 1.0f) * p.lambda_));
                // adv_t depends on v_{t+1} through the q bootstrap: weight 1 at the last step,
                // (1 - lambda) elsewhere (the lambda * vs_{t+1} part is a stop-gradient target)
                const float w = (t == T - 2) ? 1.f : p.one_minus_lambda;
                s_nx[i] = detach_adv ? 0.f : -st.lp * c_pg * u_disc * w * p.inv_n_pg;
            }
        }
        fetch_small(j + G);  // next unit's small fields: in flight during the rest of this unit
        group_sync(bar_id, gthreads);
        if (first) K1_STAMP(2);

        // ---- backward-in-time recursion e_k = td_k + (discount_k c_k) e_{k+1} ------------------
        // one register chain per column, steps fetched 6 at a time, in the reference loop's operation order
        if (i < ucols) {
            const float2 *__restrict__ qp = s_tg + i * T + (T - 2);
            float *__restrict__ ep = s_e + i * T + (T - 2);
            float e = 0.f;
            int k = T - 2;
            for (; k >= 0; k -= 6, qp -= 6, ep -= 6) {  // 6 independent loads, then 6 chained FMAs
                float2 q[6];
#pragma unroll
                for (int u = 0; u < 6; ++u) q[u] = (k - u >= 0) ? qp[-u] : make_float2(0.f, 0.f);
#pragma unroll
                for (int u = 0; u < 6; ++u)
                    if (k - u >= 0) { e = q[u].x + q[u].y * e; ep[-u] = e; }
            }
        }
        group_sync(bar_id, gthreads);
        if (first) K1_STAMP(3);

        // ---- loss terms (they need only the statistics and the recursion), per-sample gradients of lp and v ----
        float g_lp = 0.f, dv = 0.f;
        if (valid) {
            if (has_next) {
                const float e = s_e[i];
                const float tgt = e + u_v;     // stop-gradient target vs_t
                const float err = tgt - u_v;   // rlego "errors"
                float qb = u_vn;
                if (t < T - 2) {
                    const float e1 = s_e[i + 1];
                    const float tgt1 = ((e1 + u_vn) - u_vn) + u_vn;
                    // two rounded products and a rounded sum, as the reference's tensor expression evaluates it: never contracted
                    // into an FMA (which of the two products the compiler fuses differed between kernels -- tests/test_gpu_fuzz.py)
                    qb = __fadd_rn(__fmul_rn(p.lambda_, tgt1), __fmul_rn(p.one_minus_lambda, u_vn));
                }
                const float q = u_rew + u_disc * qb;
                const float qmv = q - u_v;
                const float adv = c_pg * qmv;
                sums[0] += st.lp * adv;
                sums[1] += err * err;
                float dpg_dlp = adv;  // d(lp*adv)/dlp, adv's own dependence on lp via rho below
                if (!detach_adv && !detach_rho && st.rho <= p.clip_pg_rho) dpg_dlp += st.lp * st.rho * qmv;
                g_lp = -dpg_dlp * p.inv_n_pg;
                dv = __fmul_rn(-2.f * err, p.inv_n_pg);  // explicit roundings: both K1 kernels give the same bits
                if (!detach_adv) dv = __fmaf_rn(st.lp * c_pg, p.inv_n_pg, dv);
            }
            if (t > 0) dv += s_nx[i - 1];
            sums[2] += st.H;
            sums[3] += st.kl;
            sums[4] += st.rho;
        }
        // The scalars are complete for this thread once its LAST unit is past the recursion: the CTA's reduction and
        // the publication to the collector start here and travel while the gradients below are still being written.
        if (!late_wait && !(p.debug & 8) && j + G >= n_units) { cta_reduce_publish(); reduced = true; }

        // ---- gradients ---------------------------------------------------------------------------
        if (valid) {
            p.dvalues[n0 + i] = dv;
            write_dlogits<AT, EXACT>(st, sx + i * A, A, g_lp, p.entropy_coeff * p.inv_n_all);
        }
        fence_proxy_async_smem();  // gradient rows -> visible to the bulk store
        group_sync(bar_id, gthreads);
        if (i == 0) {
            __threadfence_block();
            if (atomicAdd(&done[st_i], 1) == chunk_units(c) - 1) {  // this group finished the chunk last: send it out
                __threadfence_block();
                done[st_i] = 0;
                fence_proxy_async_smem();
                bulk_s2g(p.dlogits + (size_t)chunk_col0(c) * T * A, smem + (size_t)st_i * 2 * chunk_words,
                         (uint32_t)(chunk_cols(c) * T * A * 4));
                bulk_commit();
                if (c + NSTAGE < n_chunks) {  // refill the stage once the store has read it
                    bulk_wait_read0();
                    issue_chunk(c + NSTAGE);
                    __threadfence_block();
                    *reinterpret_cast<volatile int *>(issued + st_i) = round + 1;
                }
            }
        }
        if (first) K1_STAMP(4);
        first = false;
    }

    // ---- scalars: warps that had no unit (or INPUTS_STABLE: the workspace is the previous launch's until it is
    // complete) reduce here; then the collector warp, which parked its duty until its own gradients were out
    if (!reduced) cta_reduce_publish();
    if (is_collector) {
        __syncwarp();
        float cta_tot[5];
#pragma unroll
        for (int k = 0; k < 5; ++k) cta_tot[k] = red[k];
        const float inv_n_pg = p.inv_n_pg, inv_n_all = p.inv_n_all, ec = p.entropy_coeff;
        float *scalars = p.scalars;
        if (!(p.debug & 2)) warp_collect<5>(cta_tot, p.workspace, gen, [=](const float(&s)[5]) {
            const float pg = s[0] * inv_n_pg, td = s[1] * inv_n_pg, ent = s[2] * inv_n_all;
            scalars[B200RL_S_LOSS] = -pg + td - ec * ent;
            scalars[B200RL_S_ENTROPY] = ent;
            scalars[B200RL_S_TD] = td;
            scalars[B200RL_S_PG] = pg;
            scalars[B200RL_S_KL] = s[3] * inv_n_all;
            scalars[B200RL_S_RATIO] = s[4] * inv_n_all;
            scalars[6] = 0.f;
            scalars[7] = 0.f;
        });
    }
    K1_STAMP(5);
    if (i == 0) bulk_wait_read0();  // shared memory must outlive the bulk stores' reads
    K1_STAMP(6);
    if (TRACE && threadIdx.x == 0) ctr[4] = globaltimer_ns();
#undef K1_STAMP
}

// ------------------------------------------------------------------------ host side
static int sm_count() {
    static int cache[kMaxDevices];
    int dev = 0;
    cudaGetDevice(&dev);
    dev = (dev >= 0 && dev < kMaxDevices) ? dev : 0;
    if (!cache[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cache[dev] = n;
    }
    return cache[dev];
}

template <int AT, bool EXACT, bool TRACE, int TT, int CW, int CC, int CG>
static int32_t launch_k1_t(const K1Params &p, int grid, int threads, size_t smem, cudaStream_t stream) {
    static size_t smem_cache[kMaxDevices];
    if (int32_t rc = ensure_dynamic_smem(impala_loss_kernel<AT, EXACT, TRACE, TT, CW, CC, CG>, smem, smem_cache, "impala_loss")) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // PDL: see pdl_wait() in the kernel
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = env_int("B200RL_PDL", 1) ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&cfg, impala_loss_kernel<AT, EXACT, TRACE, TT, CW, CC, CG>, p);
    if (e != cudaSuccess) {
        set_error("impala_loss_kernel launch: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    return check_launch("impala_loss_kernel");
}

template <int AT, bool EXACT>
static int32_t launch_k1(const ImpalaArgs &a, cudaStream_t stream) {
    const int T = a.T, A = a.A, B = a.B;
    const int max_threads = k1_max_threads<AT>();
    const int max_w = max_threads / 32;
    if (T > 32 * max_w) return 1;
    // TMA bulk copies: 16-byte aligned addresses and sizes for every chunk (a whole number of units, or the
    // tail of the batch)
    if (((reinterpret_cast<uintptr_t>(a.logits) | reinterpret_cast<uintptr_t>(a.pi_ref) |
          reinterpret_cast<uintptr_t>(a.dlogits)) % 16) != 0 || ((long long)B * T) % 4 != 0)
        return 1;
    // group shape: the fewest warps whose lanes are >= 90 % samples, else the best fill
    int W = 0, C = 0;
    const int force_w = env_int("B200RL_K1_W", 0);
    double best = 0.0;
    for (int w = 1; w <= max_w; ++w) {
        int c = (32 * w) / T;
        if (c < 1) continue;
        if (c > B) c = B;
        while (c > 1 && (c * T) % 4 != 0) --c;  // units start 16-byte aligned
        if ((c * T) % 4 != 0) continue;
        const double eff = (double)c * T / (32.0 * w);
        if (force_w ? (w == force_w) : (eff > best + 1e-9)) { best = eff; W = w; C = c; }
        if (force_w ? (w == force_w) : (eff >= 0.9)) break;
    }
    if (!W) return 1;
    K1Params p;
    p.logits = a.logits; p.pi_ref = a.pi_ref; p.rewards = a.rewards; p.discounts = a.discounts; p.values = a.values;
    p.actions = a.actions; p.dlogits = a.dlogits; p.dvalues = a.dvalues; p.scalars = a.scalars; p.workspace = a.workspace;
    p.B = B; p.T = T; p.A = A;
    p.lambda_ = a.lambda_; p.one_minus_lambda = a.one_minus_lambda; p.clip_rho = a.clip_rho; p.clip_pg_rho = a.clip_pg_rho;
    p.entropy_coeff = a.entropy_coeff; p.inv_n_pg = a.inv_n_pg; p.inv_n_all = a.inv_n_all; p.flags = a.flags;
    p.C = C; p.W = W;
    p.U = (B + C - 1) / C;
    p.inv_T = (unsigned int)(((1ull << 32) + (unsigned)T - 1) / (unsigned)T);
    p.ns4 = (C * T + 3) & ~3;
    p.unit_words = C * T * A;
    // one CTA per SM (small problems: one unit per CTA first); groups per CTA: fill the CTA, at most 15 named barriers
    const int cps = max(1, min(4, env_int("B200RL_K1_CTAS_PER_SM", 1)));   // CTAs per SM the grid is sized for
    const int grid = min(sm_count() * cps, p.U);
    if (grid > kMaxPartialCtas) return 1;
    const int units_per_cta = (p.U + grid - 1) / grid;
    int G = max_threads / cps / (32 * W);
    if (W > 1 && G > 15) G = 15;
    G = max(1, min(G, env_int("B200RL_K1_G", G)));
    G = min(G, units_per_cta);
    // chunks: GPC consecutive units per pair of bulk copies; every chunk of the CTA resident when shared memory allows
    const size_t unit_bytes = (size_t)p.unit_words * 4, budget = (size_t)(200 * 1024) / cps;
    int GPC = max(1, min(env_int("B200RL_K1_GPC", 2), units_per_cta));
    while (GPC > 1 && 2 * unit_bytes * GPC * 2 > budget) --GPC;
    if (2 * unit_bytes * GPC > budget) return 1;
    const int chunks_per_cta = (units_per_cta + GPC - 1) / GPC;
    int NSTAGE = (int)min((size_t)chunks_per_cta, budget / (2 * unit_bytes * GPC));
    const int min_stages = (G + GPC - 1) / GPC + 1;  // a round of groups plus one chunk in flight
    NSTAGE = max(1, min(NSTAGE, env_int("B200RL_K1_STAGES", NSTAGE)));
    if (NSTAGE < chunks_per_cta && NSTAGE < min_stages) {  // long rollouts: fewer groups so that the ring still pipelines
        G = max(1, min(G, (NSTAGE - 1) * GPC));
    }
    p.G = G; p.GPC = GPC; p.NSTAGE = NSTAGE;
    auto magic = [](int d) { return (unsigned int)(((1ull << 32) + (unsigned)d - 1) / (unsigned)d); };
    p.inv_W = magic(W); p.inv_GPC = magic(GPC); p.inv_NSTAGE = magic(NSTAGE);
    p.units_lo = p.U / grid; p.units_rem = p.U % grid;
    p.debug = env_int("B200RL_K1_DEBUG", 0);
    p.prefetch = env_int("B200RL_K1_PREFETCH", 1);
    const int threads = G * W * 32;
    const size_t smem = (size_t)NSTAGE * 2 * GPC * unit_bytes + (size_t)G * 4 * p.ns4 * 4 + (size_t)NSTAGE * 16 +
                        5 * (size_t)(threads / 32) * 4 + 16 + 128;
    // the reference's shape class (T = 20 rollouts, procgen's 15 actions): compile-time T / W / C / GPC
    if constexpr (AT == 15 && EXACT) {
        if (T == 20 && W == 2 && C == 3 && GPC == 2 && env_int("B200RL_K1_GENERIC", 0) == 0) {
            if (a.flags & B200RL_TRACE) return launch_k1_t<AT, EXACT, true, 20, 2, 3, 2>(p, grid, threads, smem, stream);
            return launch_k1_t<AT, EXACT, false, 20, 2, 3, 2>(p, grid, threads, smem, stream);
        }
    }
    if (a.flags & B200RL_TRACE) return launch_k1_t<AT, EXACT, true, 0, 0, 0, 0>(p, grid, threads, smem, stream);
    return launch_k1_t<AT, EXACT, false, 0, 0, 0, 0>(p, grid, threads, smem, stream);
}

int32_t launch_impala_persistent(const ImpalaArgs &a, cudaStream_t stream) {
    if (a.time_major) return 1;
    switch (a.A) {  // exact-A instantiations (no padding predicates) for the common action spaces
        case 4: return launch_k1<4, true>(a, stream);
        case 6: return launch_k1<6, true>(a, stream);
        case 15: return launch_k1<15, true>(a, stream);
        case 18: return launch_k1<18, true>(a, stream);
        default: break;
    }
    if (a.A <= 8) return launch_k1<8, false>(a, stream);
    if (a.A <= 16) return launch_k1<16, false>(a, stream);
    return launch_k1<32, false>(a, stream);
}

}  // namespace b200rl

extern "C" int32_t b200rl_vtrace_impala_loss_fwdbwd(
    const float *logits, const float *pi_ref, const int64_t *actions, const float *rewards,
    const float *discounts, const float *values, int64_t B, int64_t T, int64_t A, int64_t stride_b,
    int64_t stride_t, int64_t n_global_b, double lambda_, float clip_rho, float clip_pg_rho,
    float entropy_coeff, uint32_t flags, float *dlogits, float *dvalues, float *scalars,
    void *workspace, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(logits && pi_ref && actions && rewards && discounts && values && dlogits &&
                       dvalues && scalars && workspace, -1, "impala_loss: null pointer argument");
    B200RL_REQUIRE(B >= 1 && T >= 2 && A >= 1, -2, "impala_loss: need B>=1, T>=2, A>=1 (got B=%lld T=%lld A=%lld)",
                   (long long)B, (long long)T, (long long)A);
    B200RL_REQUIRE(A <= 32, -3, "impala_loss: A=%lld > 32 actions not supported", (long long)A);
    B200RL_REQUIRE(B * T < (1ll << 31), -3, "impala_loss: B*T too large");
    const bool batch_major = (stride_t == 1 && stride_b == T);
    const bool time_major = (stride_b == 1 && stride_t == B);
    B200RL_REQUIRE(batch_major || time_major, -4,
                   "impala_loss: strides (%lld,%lld) are neither batch-major (T,1) nor time-major (1,B)",
                   (long long)stride_b, (long long)stride_t);
    B200RL_REQUIRE(n_global_b >= B, -2, "impala_loss: n_global_b < B");

    ImpalaArgs a;
    a.logits = logits; a.pi_ref = pi_ref; a.rewards = rewards; a.discounts = discounts; a.values = values;
    a.actions = reinterpret_cast<const long long *>(actions);
    a.dlogits = dlogits; a.dvalues = dvalues; a.scalars = scalars; a.workspace = workspace;
    a.B = (int)B; a.T = (int)T; a.A = (int)A;
    a.time_major = (time_major && !batch_major) ? 1 : 0;
    a.lambda_ = (float)lambda_; a.one_minus_lambda = (float)(1.0 - lambda_);
    a.clip_rho = clip_rho; a.clip_pg_rho = clip_pg_rho; a.entropy_coeff = entropy_coeff;
    a.inv_n_pg = (float)(1.0 / ((double)n_global_b * (double)(T - 1)));
    a.inv_n_all = (float)(1.0 / ((double)n_global_b * (double)T));
    a.flags = flags;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    // Two implementations, chosen by what the caller told us about the launch (B200RL_K1_TILE: 0 = this rule,
    // 1 = always the tile kernel, 2 = always the persistent kernel):
    //  * a dependent launch (the default: one learner step = one K1 between the head GEMM and its backward) has the GPU
    //    to itself, so latency is everything: the persistent, warp-group pipelined kernel (6.0-6.8 us against 8.0 us);
    //  * B200RL_INPUTS_STABLE says launches on independent rollout batches are in flight together: throughput is
    //    everything and many small CTAs of different launches interleave best: the tile kernel (2.7 us/step on 3+ streams).
    const int impl = env_int("B200RL_K1_TILE", 0);
    const bool want_persistent = impl == 2 || (impl == 0 && !(flags & B200RL_INPUTS_STABLE));
    if (!a.time_major && want_persistent) {
        const int32_t rc = launch_impala_persistent(a, s);
        if (rc != 1) return rc;
    }
    return launch_impala_tile(a, s);
}
