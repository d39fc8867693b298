// K1, tile variant: fused V-trace + IMPALA loss, forward and backward in ONE launch, one CTA per tile.
// Replaces reference agents/impala/learning.py:146-160,167-168 (+ rlego V-trace).  Used for the time-major
// layout and for shapes the persistent kernel (impala_loss.cu) does not take; B200RL_K1_IMPL=tile forces it.
//
// Work split: a CTA owns a tile of `C` batch columns x all T steps; one thread per
// (column, step) sample.  The tile's logits and behaviour logits are staged in
// shared memory by TMA 1-D bulk copies (batch-major: one contiguous copy per tensor;
// time-major: one 2-D tensor-map copy), each thread keeps its A logits in registers, the
// T-1 step V-trace recursion runs as a register chain in C threads (reference
// operation order), and d loss/d logits is written back over the staged tile and
// leaves by a bulk store.  Every input byte is read once, every gradient byte
// written once: 204 B/sample at A=15.
#include "impala_loss.cuh"
#include "softmax.cuh"

namespace b200rl {

// Tensor maps of the time-major instantiation ([T rows][B*A] tensors, box = T x C*A); a separate kernel
// parameter so that batch-major launches do not carry 384 bytes of unused maps.
template <bool TM>
struct TileMaps {};
template <>
struct TileMaps<true> {
    CUtensorMap logits, pi_ref, dlogits;
};

struct ImpalaParams {
    int use_tmap;
    const float *logits, *pi_ref, *rewards, *discounts, *values;
    const long long *actions;
    float *dlogits, *dvalues, *scalars;
    void *workspace;
    int B, T, A, C, time_major;
    int bulk_full, bulk_last;  // tiles that can use TMA bulk copies (alignment), see the launcher
    unsigned int inv_div;  // ceil(2^32 / T) batch-major, ceil(2^32 / C) time-major
    float lambda_, one_minus_lambda, clip_rho, clip_pg_rho, entropy_coeff, inv_n_pg, inv_n_all;
    uint32_t flags;
};

// TM: time-major [T,B,...] layout.  A template parameter so that the batch-major kernel carries no stride
// arithmetic and no per-row copy loop (it executes ~10% fewer instructions than a run-time switch).
template <int AT, bool EXACT, bool TRACE, bool TM>
__global__ void __launch_bounds__(AT <= 16 ? 1024 : 256) impala_loss_tile_kernel(const __grid_constant__ ImpalaParams p, const __grid_constant__ TileMaps<TM> maps) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int T = p.T, A = EXACT ? AT : p.A, C = p.C, NS = C * T;
    float *sx = reinterpret_cast<float *>(smem_raw);    // [NS*A] logits -> dlogits
    const int tile_elems = TM ? ((NS * A + 31) & ~31) : NS * A;  // tensor-map tiles start 128-byte aligned
    float *sy = sx + tile_elems;                        // [NS*A] behaviour logits
    float2 *s_tg = reinterpret_cast<float2 *>(sy + tile_elems);  // [NS] (clipped TD error, discount * c)
    float *s_e = reinterpret_cast<float *>(s_tg + NS);  // [NS] V-trace errors e_t
    float *s_nx = s_e + NS;                             // [NS] d loss/d v_{t+1} pushed by step t
    float *red = s_nx + NS;                             // [5*blockDim]
    uint64_t *bar = reinterpret_cast<uint64_t *>(red + 5 * blockDim.x);

    // B200RL_TRACE selects a separate instantiation: the production kernel carries no stamp code
    const bool trace = TRACE && threadIdx.x == 0 && blockIdx.x < kMaxTraceCtas;
    unsigned long long *tr = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(p.workspace) + kTraceOffset) + blockIdx.x * 16;
#define K1_STAMP(i) do { if (TRACE && trace) { tr[i] = globaltimer_ns(); tr[8 + i] = (unsigned long long)clock64(); } } while (0)
    K1_STAMP(0);
    if (trace) { unsigned int smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid)); tr[7] = smid; }
    pdl_launch_dependents();
    const bool late_wait = p.flags & B200RL_INPUTS_STABLE;
    if (!late_wait) pdl_wait();  // inputs may come from the preceding kernel

    const int b0 = blockIdx.x * C;
    const int cols = min(C, p.B - b0);
    const int s = threadIdx.x;
    int c, t;
    // s / T (resp. s / C) by one IMAD.HI with the host-computed ceil(2^32 / divisor): exact for s * divisor < 2^32
    if (TM) { t = C == 1 ? s : (int)__umulhi((unsigned)s, p.inv_div); c = s - t * C; }
    else { c = (int)__umulhi((unsigned)s, p.inv_div); t = s - c * T; }
    const bool valid = s < NS && c < cols;
    const int st_local = TM ? C : 1;  // local index distance between t and t+1
    const int n = TM ? t * p.B + b0 + c : (b0 + c) * T + t;
    const int n_stride_t = TM ? p.B : 1;

    // ---- phase 1: stage the tile ------------------------------------------------------
    const int row_elems = (TM ? cols : cols * T) * A;  // contiguous run
    const int nrows = TM ? T : 1;
    const size_t g_row0 = (size_t)(TM ? b0 : b0 * T) * A;
    const size_t g_row_stride = (size_t)p.B * A;  // time-major only
    const int s_row_stride = C * A;               // time-major only
    const uint32_t row_bytes = (uint32_t)row_elems * 4u;
    // TMA bulk copies need 16-byte aligned addresses and sizes: decided on the host (full tiles / the ragged
    // last tile), since every term depends on launch arguments only
    // time-major with a tensor map: ONE 2-D tiled copy per tensor (T rows x C*A floats; the ragged last tile
    // is zero-filled / clipped by the TMA unit) instead of T row copies
    const bool tmap = TM && p.use_tmap;
    const bool bulk_ok = tmap || ((cols == C) ? p.bulk_full : p.bulk_last);
    // The copies are issued by lane 0 of every warp, rows dealt round-robin: a UBLKCP takes tens of
    // cycles to issue, so one thread issuing 2*T of them (time-major) would delay the last row by
    // microseconds.  Batch-major has one row per tensor: thread 0 alone, no extra barrier.
    const int warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const bool issuer = (TM && !tmap) ? ((threadIdx.x & 31) == 0 && warp < nrows) : threadIdx.x == 0;
    if (bulk_ok) {
        if (threadIdx.x == 0) {
            mbar_init(bar, 1);
            fence_mbar_init();
            mbar_expect_tx(bar, tmap ? 2u * (uint32_t)(NS * A) * 4u : 2u * row_bytes * nrows);
        }
        if constexpr (TM) {
            if (tmap && threadIdx.x == 0) {
                tma_load_2d(sx, &maps.logits, b0 * A, 0, bar);
                tma_load_2d(sy, &maps.pi_ref, b0 * A, 0, bar);
            }
        }
        if (tmap) {
        } else if (TM) {
            __syncthreads();  // barrier initialised before other warps' copies can complete on it
            if (issuer) {
                for (int r = warp; r < nrows; r += nwarps) {
                    bulk_g2s(sx + r * s_row_stride, p.logits + g_row0 + r * g_row_stride, row_bytes, bar);
                    bulk_g2s(sy + r * s_row_stride, p.pi_ref + g_row0 + r * g_row_stride, row_bytes, bar);
                }
            }
        } else if (threadIdx.x == 0) {
            bulk_g2s(sx, p.logits + g_row0, row_bytes, bar);
            bulk_g2s(sy, p.pi_ref + g_row0, row_bytes, bar);
        }
    } else {
        for (int r = 0; r < nrows; ++r)
            for (int i = threadIdx.x; i < row_elems; i += blockDim.x) {
                sx[r * s_row_stride + i] = ld_stream_f32(p.logits + g_row0 + r * g_row_stride + i);
                sy[r * s_row_stride + i] = ld_stream_f32(p.pi_ref + g_row0 + r * g_row_stride + i);
            }
    }
    long long act = 0;
    float rew = 0.f, disc = 0.f, v = 0.f, v_next = 0.f;
    const bool has_next = valid && t < T - 1;
    if (valid) {
        act = __ldg(p.actions + n);
        rew = __ldg(p.rewards + n);
        disc = __ldg(p.discounts + n);
        v = __ldg(p.values + n);
        if (has_next) v_next = __ldg(p.values + n + n_stride_t);
    }
    __syncthreads();  // mbarrier init visible / cooperative stores done
    if (bulk_ok) mbar_wait(bar, 0);
    K1_STAMP(1);

    // ---- phase 2: per-sample softmax statistics, TD error, trace coefficient -----------
    SampleStats<AT> st;
    float c_pg = 0.f;
    const bool detach_adv = p.flags & B200RL_DETACH_ADV, detach_rho = p.flags & B200RL_DETACH_RHO;
    if (valid) {
        sample_stats<AT, EXACT>(st, sx + s * A, sy + s * A, A, act);
        c_pg = fminf(st.rho, p.clip_pg_rho);
        if (has_next) {
            s_tg[s] = make_float2(fminf(st.rho, p.clip_rho) * (rew + disc * v_next - v),
                                  disc * (fminf(st.rho, 1.0f) * p.lambda_));
            // adv_t depends on v_{t+1} through the q bootstrap: weight 1 at the last step,
            // (1 - lambda) elsewhere (the lambda * vs_{t+1} part is a stop-gradient target)
            const float w = (t == T - 2) ? 1.f : p.one_minus_lambda;
            s_nx[s] = detach_adv ? 0.f : -st.lp * c_pg * disc * w * p.inv_n_pg;
        }
    }
    __syncthreads();
    K1_STAMP(2);

    // ---- phase 3: backward-in-time recursion e_k = td_k + (discount_k c_k) e_{k+1} ----------
    // One register chain per column (the column's first thread), steps fetched 4 at a time, in the
    // reference loop's operation order; results go to s_e.
    float e_t = 0.f, e_t1 = 0.f;  // e_t, e_{t+1} of this thread's sample
    if (s < cols) {
        const int base = TM ? s : s * T;
        const float2 *__restrict__ qp = s_tg + base + (T - 2) * st_local;  // step k
        float *__restrict__ ep = s_e + base + (T - 2) * st_local;
        const int st4 = 4 * st_local;
        float e = 0.f;
        int k = T - 2;
        for (; k >= 3; k -= 4, qp -= st4, ep -= st4) {  // 4 independent loads, then 4 chained FMAs
            const float2 q0 = qp[0], q1 = qp[-st_local], q2 = qp[-2 * st_local], q3 = qp[-3 * st_local];
            e = q0.x + q0.y * e; ep[0] = e;
            e = q1.x + q1.y * e; ep[-st_local] = e;
            e = q2.x + q2.y * e; ep[-2 * st_local] = e;
            e = q3.x + q3.y * e; ep[-3 * st_local] = e;
        }
        for (; k >= 0; --k, qp -= st_local, ep -= st_local) {
            const float2 q = *qp;
            e = q.x + q.y * e;
            *ep = e;
        }
    }
    __syncthreads();
    if (has_next) {
        e_t = s_e[s];
        if (t < T - 2) e_t1 = s_e[s + st_local];
    }
    K1_STAMP(3);

    // ---- phase 4: losses and gradients -------------------------------------------------
    float sums[5] = {0.f, 0.f, 0.f, 0.f, 0.f};  // pg, value, entropy, kl, ratio
    if (valid) {
        float g_lp = 0.f, dv = 0.f;
        if (has_next) {
            const float e = e_t;
            const float tgt = e + v;     // stop-gradient target vs_t
            const float err = tgt - v;   // rlego "errors"
            float qb = v_next;
            if (t < T - 2) {
                const float e1 = e_t1;
                const float tgt1 = ((e1 + v_next) - v_next) + v_next;
                // two rounded products and a rounded sum, as the reference's tensor expression evaluates it: never contracted
                // into an FMA (which of the two products the compiler fuses differed between kernels -- tests/test_gpu_fuzz.py)
                qb = __fadd_rn(__fmul_rn(p.lambda_, tgt1), __fmul_rn(p.one_minus_lambda, v_next));
            }
            const float q = rew + disc * qb;
            const float qmv = q - v;
            const float adv = c_pg * qmv;
            sums[0] = st.lp * adv;
            sums[1] = err * err;
            float dpg_dlp = adv;  // d(lp*adv)/dlp, adv's own dependence on lp via rho below
            if (!detach_adv && !detach_rho && st.rho <= p.clip_pg_rho) dpg_dlp += st.lp * st.rho * qmv;
            g_lp = -dpg_dlp * p.inv_n_pg;
            dv = __fmul_rn(-2.f * err, p.inv_n_pg);  // explicit roundings: both K1 kernels give the same bits
            if (!detach_adv) dv = __fmaf_rn(st.lp * c_pg, p.inv_n_pg, dv);
        }
        if (t > 0) dv += s_nx[s - st_local];
        p.dvalues[n] = dv;
        sums[2] = st.H;
        sums[3] = st.kl;
        sums[4] = st.rho;
        write_dlogits<AT, EXACT>(st, sx + s * A, A, g_lp, p.entropy_coeff * p.inv_n_all);
    }
    if (bulk_ok) fence_proxy_async_smem();
    __syncthreads();
    K1_STAMP(4);
    if (bulk_ok) {
        if (issuer) {
            if (tmap) {
                if constexpr (TM) tma_store_2d(&maps.dlogits, sx, b0 * A, 0);
            } else if (TM) {
                for (int r = warp; r < nrows; r += nwarps)
                    bulk_s2g(p.dlogits + g_row0 + r * g_row_stride, sx + r * s_row_stride, row_bytes);
            } else {
                bulk_s2g(p.dlogits + g_row0, sx, row_bytes);
            }
            bulk_commit();
        }
    } else {
        for (int r = 0; r < nrows; ++r)
            for (int i = threadIdx.x; i < row_elems; i += blockDim.x)
                p.dlogits[g_row0 + r * g_row_stride + i] = sx[r * s_row_stride + i];
    }

    // ---- scalars: deterministic grid reduction -----------------------------------------
    const float inv_n_pg = p.inv_n_pg, inv_n_all = p.inv_n_all, ec = p.entropy_coeff;
    float *scalars = p.scalars;
    grid_reduce<5>(
        sums, red, p.workspace,
        [=](const float(&tot)[5]) {
            const float pg = tot[0] * inv_n_pg, td = tot[1] * inv_n_pg, ent = tot[2] * inv_n_all;
            scalars[B200RL_S_LOSS] = -pg + td - ec * ent;
            scalars[B200RL_S_ENTROPY] = ent;
            scalars[B200RL_S_TD] = td;
            scalars[B200RL_S_PG] = pg;
            scalars[B200RL_S_KL] = tot[3] * inv_n_all;
            scalars[B200RL_S_RATIO] = tot[4] * inv_n_all;
            scalars[6] = 0.f;
            scalars[7] = 0.f;
        },
        [=]() { if (late_wait) pdl_wait(); });  // the workspace and scalars are shared with the previous launch
    K1_STAMP(5);
    if (bulk_ok && issuer) bulk_wait_read0();  // smem must outlive the bulk store's reads
    K1_STAMP(6);
#undef K1_STAMP
}

}  // namespace b200rl

// ------------------------------------------------------------------------ host side
namespace b200rl {

template <int AT, bool EXACT, bool TRACE, bool TM>
static int32_t launch_impala_t(ImpalaParams &p, int threads, size_t smem, cudaStream_t stream) {
    TileMaps<TM> maps;
    if constexpr (TM) {
        if (env_int("B200RL_K1_TMAP", 1)) {
            const uint64_t inner = (uint64_t)p.B * p.A, pitch = inner * 4;
            const uint32_t bi = (uint32_t)(p.C * p.A), br = (uint32_t)p.T;
            p.use_tmap = encode_tmap_2d_f32(&maps.logits, p.logits, inner, br, pitch, bi, br) &&
                         encode_tmap_2d_f32(&maps.pi_ref, p.pi_ref, inner, br, pitch, bi, br) &&
                         encode_tmap_2d_f32(&maps.dlogits, p.dlogits, inner, br, pitch, bi, br);
        }
    }
    static size_t smem_cache[kMaxDevices];
    if (int32_t rc = ensure_dynamic_smem(impala_loss_tile_kernel<AT, EXACT, TRACE, TM>, smem, smem_cache, "impala_loss")) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((p.B + p.C - 1) / p.C));
    cfg.blockDim = dim3((unsigned)threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // PDL: see pdl_wait() in the kernel
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = env_int("B200RL_PDL", 1) ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&cfg, impala_loss_tile_kernel<AT, EXACT, TRACE, TM>, p, maps);
    if (e != cudaSuccess) {
        set_error("impala_loss_tile_kernel launch: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    return check_launch("impala_loss_tile_kernel");
}

template <int AT, bool EXACT>
static int32_t launch_impala(ImpalaParams &p, int threads, size_t smem, cudaStream_t stream) {
    if (p.flags & B200RL_TRACE) {  // stamped instantiations (scripts/k1_trace.py)
        if (p.time_major) return launch_impala_t<AT, EXACT, true, true>(p, threads, smem, stream);
        return launch_impala_t<AT, EXACT, true, false>(p, threads, smem, stream);
    }
    if (p.time_major) return launch_impala_t<AT, EXACT, false, true>(p, threads, smem, stream);
    return launch_impala_t<AT, EXACT, false, false>(p, threads, smem, stream);
}

}  // namespace b200rl

namespace b200rl {

int32_t launch_impala_tile(const ImpalaArgs &a, cudaStream_t s) {
    const long long B = a.B, T = a.T, A = a.A;
    const bool time_major = a.time_major != 0, batch_major = !time_major;
    const int max_threads = A <= 16 ? 640 : 256;  // == __launch_bounds__ of the instantiation
    B200RL_REQUIRE(T <= max_threads, -3, "impala_loss: T=%lld > %d steps not supported at A=%lld",
                   (long long)T, max_threads, (long long)A);
    const float *logits = a.logits, *pi_ref = a.pi_ref;
    float *dlogits = a.dlogits;

    ImpalaParams p;
    p.logits = a.logits; p.pi_ref = a.pi_ref; p.rewards = a.rewards; p.discounts = a.discounts; p.values = a.values;
    p.actions = a.actions;
    p.dlogits = a.dlogits; p.dvalues = a.dvalues; p.scalars = a.scalars; p.workspace = a.workspace;
    p.B = a.B; p.T = a.T; p.A = a.A;
    p.time_major = a.time_major;
    p.lambda_ = a.lambda_; p.one_minus_lambda = a.one_minus_lambda; p.clip_rho = a.clip_rho; p.clip_pg_rho = a.clip_pg_rho; p.entropy_coeff = a.entropy_coeff;
    p.inv_n_pg = a.inv_n_pg;
    p.inv_n_all = a.inv_n_all;
    p.flags = a.flags;

    // columns per CTA: a multiple of 4 so time-major rows stay 16-byte aligned
    int target = env_int("B200RL_K1_SAMPLES", 160);
    int C = (int)(target / T);
    if (C >= 4 && time_major && !batch_major) C &= ~3;
    if (C < 1) C = 1;
    if (C > B) C = (int)B;
    while (C > 1 && C * T > max_threads) C = (C > 4) ? ((C - 1) & ~3) : C - 1;
    auto threads_for = [&](int c) { return ((c * (int)T + 31) / 32) * 32; };
    auto smem_for = [&](int c) { return (size_t)c * T * (2 * A + 4) * 4 + 5 * (size_t)threads_for(c) * 4 + 16 + 256; };
    while (C > 1 && smem_for(C) > 200 * 1024) C = (C > 4) ? ((C / 2) & ~3) : C - 1;
    B200RL_REQUIRE(smem_for(C) <= 226 * 1024, -3, "impala_loss: T*A too large for one CTA tile");
    if ((B + C - 1) / C > kMaxPartialCtas) {
        // very wide batches of long rollouts: the grid may not exceed the partials the workspace holds, so a CTA takes
        // more columns than the tuned tile (found by tests/test_gpu_fuzz.py: T=128 used to stop at B=16384)
        int need = (int)((B + kMaxPartialCtas - 1) / kMaxPartialCtas);
        if (need >= 4 && time_major && !batch_major) need = (need + 3) & ~3;
        B200RL_REQUIRE((long long)need * T <= max_threads && smem_for(need) <= 226 * 1024, -3,
                       "impala_loss: B=%lld columns of T=%lld need %d columns per CTA to fit the %d partial sums of the workspace, "
                       "which one CTA cannot hold -- split the batch", (long long)B, (long long)T, need, kMaxPartialCtas);
        C = need;
    }
    p.C = C;
    {   // alignment of the bulk copies: row start, row size, row strides (time-major), base pointers
        const bool tm = p.time_major != 0;
        const bool ptrs = ((reinterpret_cast<uintptr_t>(logits) | reinterpret_cast<uintptr_t>(pi_ref) |
                            reinterpret_cast<uintptr_t>(dlogits)) % 16) == 0;
        auto ok = [&](long long cols) {
            const long long row_bytes = (tm ? cols : cols * T) * A * 4;
            const long long tile_step = (tm ? (long long)C : (long long)C * T) * A * 4;  // distance between tiles' first rows
            bool a = ptrs && row_bytes % 16 == 0 && tile_step % 16 == 0;
            if (tm) a = a && ((long long)B * A * 4) % 16 == 0 && ((long long)C * A * 4) % 16 == 0;
            return a ? 1 : 0;
        };
        p.bulk_full = ok(C);
        p.bulk_last = (B % C) ? ok(B % C) : p.bulk_full;
    }
    p.use_tmap = 0;  // set by the time-major launcher
    p.inv_div = (unsigned int)(((1ull << 32) + (unsigned)(p.time_major ? C : (int)T) - 1) / (unsigned)(p.time_major ? C : (int)T));
    const int threads = threads_for(C);
    const size_t smem = smem_for(C);
    switch (A) {  // exact-A instantiations (no padding predicates) for the common action spaces
        case 4: return launch_impala<4, true>(p, threads, smem, s);
        case 6: return launch_impala<6, true>(p, threads, smem, s);
        case 15: return launch_impala<15, true>(p, threads, smem, s);
        case 18: return launch_impala<18, true>(p, threads, smem, s);
        default: break;
    }
    if (A <= 8) return launch_impala<8, false>(p, threads, smem, s);
    if (A <= 16) return launch_impala<16, false>(p, threads, smem, s);
    return launch_impala<32, false>(p, threads, smem, s);
}

}  // namespace b200rl
