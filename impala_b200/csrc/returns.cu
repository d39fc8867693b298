// K2: lambda-returns and GAE time scans, batched over columns.
// Replaces rlego.lambda_returns (reference agents/ppo/learning.py:65-69) and
// rlego.truncated_generalized_advantage_estimation (losses.py:11).
//
// Both are the backward affine recurrence  y_t = b_t + a_t * y_{t+1}:
//   lambda-returns: a_t = g_t*l,  b_t = r_t + g_t*(1-l)*v_t[t],       y_T = v_t[T-1]
//   GAE:            a_t = l*g_t,  b_t = r_t + g_t*V[t+1] - V[t],      y_T = 0
//
// Time-major kernel (HBM-bound, 16 B/sample): a CTA owns 32 adjacent columns (lane =
// column, every load a full 128-byte line) and splits T into chunks of L steps, one
// warp per chunk.  Pass 1: each warp loads its L x 3 values into registers (3L
// independent loads in flight per thread) and composes its chunk's affine map.  The
// maps meet in shared memory, each warp folds the later chunks into its carry-in
// (<= 31 FMAs), then pass 2 re-walks the chunk from registers in the reference's
// operation order and stores.  No second trip to HBM.
#include "common.cuh"

namespace b200rl {

int env_int(const char *name, int dflt);

enum { MODE_LAMBDA_RETURNS = 0, MODE_GAE = 1 };

struct ScanParams {
    const float *r, *disc, *v;
    float *out;
    int B, T;
    long long sb, st, vsb, vst;  // element strides of r/disc/out and of v
    float lambda_, one_minus_lambda;  // both rounded to fp32 from the double argument
};

// One step in the reference's operation order with every product and sum rounded separately
// (no FMA contraction), i.e. exactly what the eager torch ops of the oracle compute; the
// kernels are HBM-bound, the extra instructions are free.
template <int MODE>
__device__ __forceinline__ float scan_step(float r, float d, float v0, float v1, float lam, float oml, float y) {
    if (MODE == MODE_LAMBDA_RETURNS)  // r + d * ((1 - lam) * v_t[t] + lam * y)
        return __fadd_rn(r, __fmul_rn(d, __fadd_rn(__fmul_rn(oml, v0), __fmul_rn(lam, y))));
    // (r + d * V[t+1] - V[t]) + lam * d * y
    return __fadd_rn(__fsub_rn(__fadd_rn(r, __fmul_rn(d, v1)), v0), __fmul_rn(__fmul_rn(lam, d), y));
}

template <int MODE, int L>
__global__ void __launch_bounds__(512) scan_time_major_kernel(const ScanParams p) {
    __shared__ float sA[16][33], sB[16][33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, K = blockDim.x >> 5;
    const int b = blockIdx.x * 32 + lane;
    const bool col_ok = b < p.B;
    const int t0 = w * L;
    float r[L], d[L], v[L + 1];
#pragma unroll
    for (int k = 0; k < L; ++k) {
        const int t = t0 + k;
        const bool ok = col_ok && t < p.T;
        const long long i = (long long)t * p.st + (long long)b * p.sb;
        r[k] = ok ? __ldg(p.r + i) : 0.f;
        d[k] = ok ? __ldg(p.disc + i) : 0.f;
        const bool vok = col_ok && t < p.T + (MODE == MODE_GAE ? 1 : 0);  // GAE: V has T+1 steps
        v[k] = vok ? __ldg(p.v + (long long)t * p.vst + (long long)b * p.vsb) : 0.f;
    }
    if (MODE == MODE_GAE) {
        const int t = t0 + L;  // V[t+1] of the chunk's last step
        v[L] = (col_ok && t <= p.T) ? __ldg(p.v + (long long)t * p.vst + (long long)b * p.vsb) : 0.f;
    } else {
        v[L] = 0.f;
    }
    // pass 1: compose the chunk's map y_in -> y_out = Bc + Ac * y_in (steps beyond T are identity)
    float Ac = 1.f, Bc = 0.f;
#pragma unroll
    for (int k = L - 1; k >= 0; --k) {
        if (t0 + k < p.T) {
            const float a = p.lambda_ * d[k];
            const float bb = scan_step<MODE>(r[k], d[k], v[k], v[k + 1], p.lambda_, p.one_minus_lambda, 0.f);
            Bc = bb + a * Bc;
            Ac = a * Ac;
        }
    }
    sA[w][lane] = Ac;
    sB[w][lane] = Bc;
    __syncthreads();
    // carry-in: fold the later chunks, starting from the bootstrap
    float y = 0.f;
    if (MODE == MODE_LAMBDA_RETURNS && col_ok)
        y = __ldg(p.v + (long long)(p.T - 1) * p.vst + (long long)b * p.vsb);
    for (int k = K - 1; k > w; --k) y = sB[k][lane] + sA[k][lane] * y;
    // pass 2: the chunk itself, reference operation order, from registers
#pragma unroll
    for (int k = L - 1; k >= 0; --k) {
        const int t = t0 + k;
        if (col_ok && t < p.T) {
            y = scan_step<MODE>(r[k], d[k], v[k], v[k + 1], p.lambda_, p.one_minus_lambda, y);
            p.out[(long long)t * p.st + (long long)b * p.sb] = y;
        }
    }
}

// Time-major kernel, shared-memory staged (default when the tile fits: T <= kSmemScanMaxT).
// A CTA owns 32 adjacent columns x all T steps.  The three input tiles go HBM -> shared memory
// with 16-byte cp.async (LDGSTS: no registers held while the loads fly, so two CTAs fit per SM
// and one CTA's loads overlap the other's arithmetic and stores).  Then the same two passes as
// above, but reading shared memory (lane = column: conflict free): pass 1 composes each warp's
// chunk map, the maps meet in shared memory, pass 2 re-walks the chunk in the reference's
// operation order and stores 128-byte lines.
constexpr int kSmemScanMaxT = 256;

template <int MODE, int kSmemScanWarps>
__global__ void __launch_bounds__(kSmemScanWarps * 32) scan_smem_kernel(const ScanParams p) {
    extern __shared__ __align__(16) float smem_scan[];
    const int T = p.T, TV = T + (MODE == MODE_GAE ? 1 : 0);
    float *sr = smem_scan;              // [T][32]
    float *sd = sr + T * 32;            // [T][32]
    float *sv = sd + T * 32;            // [TV][32]
    float *sA = sv + TV * 32;           // [kSmemScanWarps][32]
    float *sB = sA + kSmemScanWarps * 32;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int b0 = blockIdx.x * 32;
    // The kernel that consumes the targets (the fused PPO loss, launched with the PDL attribute) may be scheduled while
    // this one runs: it waits for this grid's completion before it reads anything (griddepcontrol.wait), but its prologue
    // and its L2 prefetch of the logits overlap this kernel.  A no-op for dependents launched without the attribute.
    pdl_launch_dependents();
    // Launched with the PDL attribute itself (in the PPO pipeline it follows the previous step's loss kernel): before the
    // dependency wait it may only HINT -- one L2 prefetch per 128-byte row segment of its three inputs -- then it waits.
    for (int row = threadIdx.x; row < TV; row += kSmemScanWarps * 32) {
        const long long col = b0 < p.B ? b0 : 0;
        if (row < T) {
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p.r + (long long)row * p.st + col));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p.disc + (long long)row * p.st + col));
        }
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p.v + (long long)row * p.vst + col));
    }
    pdl_wait();
    // ---- stage: chunk = (row, group of 4 columns); rows are st (vst) elements apart in HBM
    for (int ch = threadIdx.x; ch < TV * 8; ch += kSmemScanWarps * 32) {
        const int row = ch >> 3, c4 = (ch & 7) * 4;
        const int remaining = p.B - (b0 + c4);  // columns left from this chunk's first column
        const int bytes = remaining >= 4 ? 16 : (remaining > 0 ? remaining * 4 : 0);
        const long long col = remaining > 0 ? b0 + c4 : 0;  // keep the address valid when nothing is read
        if (row < T) {
            cp_async16(sr + row * 32 + c4, p.r + (long long)row * p.st + col, bytes);
            cp_async16(sd + row * 32 + c4, p.disc + (long long)row * p.st + col, bytes);
        }
        cp_async16(sv + row * 32 + c4, p.v + (long long)row * p.vst + col, bytes);
    }
    cp_async_wait_all();
    __syncthreads();
    const int b = b0 + lane;
    const bool col_ok = b < p.B;
    const int L = (T + kSmemScanWarps - 1) / kSmemScanWarps;
    const int t0 = w * L, t1 = min(T, t0 + L);
    // pass 1: compose the chunk's map y_in -> Bc + Ac * y_in (one warp: nothing to compose, and the
    // whole column is one chain in the reference's order -- bit exact)
    float Ac = 1.f, Bc = 0.f;
    for (int t = (kSmemScanWarps > 1 ? t1 : t0) - 1; t >= t0; --t) {
        const float d = sd[t * 32 + lane];
        const float v1 = (MODE == MODE_GAE) ? sv[(t + 1) * 32 + lane] : 0.f;
        const float a = p.lambda_ * d;
        const float bb = scan_step<MODE>(sr[t * 32 + lane], d, sv[t * 32 + lane], v1, p.lambda_, p.one_minus_lambda, 0.f);
        Bc = bb + a * Bc;
        Ac = a * Ac;
    }
    sA[w * 32 + lane] = Ac;
    sB[w * 32 + lane] = Bc;
    __syncthreads();
    // carry-in: fold the later chunks, starting from the bootstrap
    float y = (MODE == MODE_LAMBDA_RETURNS) ? sv[(T - 1) * 32 + lane] : 0.f;
    for (int k = kSmemScanWarps - 1; k > w; --k) y = sB[k * 32 + lane] + sA[k * 32 + lane] * y;
    // pass 2: the chunk itself, reference operation order
    for (int t = t1 - 1; t >= t0; --t) {
        const float v1 = (MODE == MODE_GAE) ? sv[(t + 1) * 32 + lane] : 0.f;
        y = scan_step<MODE>(sr[t * 32 + lane], sd[t * 32 + lane], sv[t * 32 + lane], v1, p.lambda_, p.one_minus_lambda, y);
        if (col_ok) __stcs(p.out + (long long)t * p.st + b, y);
    }
}

// (A thread-block-cluster version -- the time axis split over the CTAs of a cluster, chunk maps exchanged through
// distributed shared memory -- was built and measured in round 1: slower than the one-CTA kernels here (T=256, B=8192:
// 6.8 vs 5.0 us; T=512, B=4096: 9.3 vs 8.7 us; co-scheduling the cluster and its two cluster barriers cost more than the
// wider rows save).  It was removed from the library in round 2; DESIGN.md keeps the numbers.)

template <int MODE>
__global__ void __launch_bounds__(128) scan_generic_kernel(const ScanParams p) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.B) return;
    const float *vb = p.v + (long long)b * p.vsb;
    float y = (MODE == MODE_LAMBDA_RETURNS) ? vb[(long long)(p.T - 1) * p.vst] : 0.f;
    float v1 = (MODE == MODE_GAE) ? vb[(long long)p.T * p.vst] : 0.f;
    for (int t = p.T - 1; t >= 0; --t) {
        const long long i = (long long)b * p.sb + (long long)t * p.st;
        const float v0 = vb[(long long)t * p.vst];
        y = scan_step<MODE>(p.r[i], p.disc[i], v0, v1, p.lambda_, p.one_minus_lambda, y);
        p.out[i] = y;
        v1 = v0;
    }
}

template <int MODE>
static int32_t launch_scan(const ScanParams &p, cudaStream_t stream) {
    const bool time_major = (p.sb == 1 && p.vsb == 1);
    const bool al16 = ((reinterpret_cast<uintptr_t>(p.r) | reinterpret_cast<uintptr_t>(p.disc) |
                        reinterpret_cast<uintptr_t>(p.v)) % 16 == 0) && p.st % 4 == 0 && p.vst % 4 == 0;
    if (time_major && al16 && p.T <= kSmemScanMaxT && p.T >= 32 && env_int("B200RL_K2_SMEM", 1)) {
        const int TV = p.T + (MODE == MODE_GAE ? 1 : 0);
        const int nw = env_int("B200RL_K2_WARPS", 8);
        const size_t smem = (size_t)(2 * p.T + TV + 2 * 8) * 32 * 4;
        static size_t c8[kMaxDevices], c4[kMaxDevices], c2[kMaxDevices], c1[kMaxDevices];
        int32_t rc = 0;
        switch (nw) {
            case 1: rc = ensure_dynamic_smem(scan_smem_kernel<MODE, 1>, smem, c1, "scan"); break;
            case 2: rc = ensure_dynamic_smem(scan_smem_kernel<MODE, 2>, smem, c2, "scan"); break;
            case 4: rc = ensure_dynamic_smem(scan_smem_kernel<MODE, 4>, smem, c4, "scan"); break;
            default: rc = ensure_dynamic_smem(scan_smem_kernel<MODE, 8>, smem, c8, "scan"); break;
        }
        if (rc) return rc;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)((p.B + 31) / 32));
        cfg.dynamicSmemBytes = smem;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // PDL: see pdl_wait() in the kernel
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        // measured A/B on one box (T=256, K2 + K3 per step): early launch + prefetch is -2 % at 8,192 columns (75.2 -> 73.6 us)
        // but +4 % at 1,024 / 2,048 (16.0 -> 16.6, 23.7 -> 24.3 us): only when the scan fills the machine (>= 1 CTA per SM)
        const int k2_pdl = env_int("B200RL_K2_PDL", -1);
        cfg.numAttrs = (env_int("B200RL_PDL", 1) && (k2_pdl == 1 || (k2_pdl < 0 && cfg.gridDim.x >= 148))) ? 1 : 0;
        cudaError_t e;
        switch (nw) {
            case 1: cfg.blockDim = dim3(32); e = cudaLaunchKernelEx(&cfg, scan_smem_kernel<MODE, 1>, p); break;
            case 2: cfg.blockDim = dim3(64); e = cudaLaunchKernelEx(&cfg, scan_smem_kernel<MODE, 2>, p); break;
            case 4: cfg.blockDim = dim3(128); e = cudaLaunchKernelEx(&cfg, scan_smem_kernel<MODE, 4>, p); break;
            default: cfg.blockDim = dim3(256); e = cudaLaunchKernelEx(&cfg, scan_smem_kernel<MODE, 8>, p); break;
        }
        if (e != cudaSuccess) {
            set_error("scan_smem_kernel launch: %s", cudaGetErrorString(e));
            (void)cudaGetLastError();
            return (int32_t)e;
        }
        return check_launch("scan_smem_kernel");
    }
    if (time_major && p.T <= 16 * 32) {  // <= 16 chunks (warps) of L steps
        const int grid = (p.B + 31) / 32;
        if (p.T <= 16 * 8) {
            const int K = (p.T + 7) / 8;
            scan_time_major_kernel<MODE, 8><<<grid, K * 32, 0, stream>>>(p);
        } else if (p.T <= 16 * 16) {
            const int K = (p.T + 15) / 16;
            scan_time_major_kernel<MODE, 16><<<grid, K * 32, 0, stream>>>(p);
        } else {
            const int K = (p.T + 31) / 32;
            scan_time_major_kernel<MODE, 32><<<grid, K * 32, 0, stream>>>(p);
        }
        return check_launch("scan_time_major_kernel");
    }
    scan_generic_kernel<MODE><<<(p.B + 127) / 128, 128, 0, stream>>>(p);
    return check_launch("scan_generic_kernel");
}

// N4: n-step bootstrapped returns (rlax / rlego n_step_bootstrapped_returns; reference
// agents/dqn/learning.py:62,79).  Every output is an independent n-step window, so one thread per
// (b,t) walks its window backwards with the lambda-returns step (same operation order as the
// oracle: bit-exact); steps past the end of the rollout are the reference's padding
// (r = 0, discount = 1, v = v_t[T-1]).  Threads are enumerated in memory order: coalesced for
// either layout; the window re-reads hit L1.
__global__ void __launch_bounds__(256) n_step_kernel(const ScanParams p, int n_step) {
    const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= (long long)p.B * p.T) return;
    int b, t;
    if (p.st == 1) { b = (int)(k / p.T); t = (int)(k - (long long)b * p.T); }
    else { t = (int)(k / p.B); b = (int)(k - (long long)t * p.B); }
    const long long base = (long long)b * p.sb;
    const float v_last = __ldg(p.v + base + (long long)(p.T - 1) * p.st);
    const int last = t + n_step - 1;
    float y = last < p.T ? __ldg(p.v + base + (long long)last * p.st) : v_last;
    for (int i = n_step - 1; i >= 0; --i) {
        const int kk = t + i;
        if (kk < p.T) {
            const long long j = base + (long long)kk * p.st;
            y = scan_step<MODE_LAMBDA_RETURNS>(__ldg(p.r + j), __ldg(p.disc + j), __ldg(p.v + j), 0.f, p.lambda_,
                                               p.one_minus_lambda, y);
        } else {
            y = scan_step<MODE_LAMBDA_RETURNS>(0.f, 1.f, v_last, 0.f, p.lambda_, p.one_minus_lambda, y);
        }
    }
    p.out[base + (long long)t * p.st] = y;
}

// N4, time-major, shared-memory staged: a CTA owns 128 columns x tile_rows consecutive steps (+ the
// n-1 steps of look-ahead), staged with 16-byte cp.async so every input byte crosses HBM once
// (+ (n-1)/tile_rows halo, usually an L2 hit).  A thread owns 4 adjacent columns: 16-byte shared-memory
// reads, one 16-byte store per output row (512 contiguous bytes per warp).  Windows do not depend on
// each other, so unlike the scans the time axis tiles freely: many small CTAs per SM, loads /
// arithmetic / stores of different CTAs overlap.
constexpr int kNStepRows = 16, kNStepWarps = 4, kNStepMaxHalo = 64, kNStepCols = 128;  // kNStepRows: default tile height (B200RL_N4_ROWS)

__global__ void __launch_bounds__(kNStepWarps * 32) n_step_smem_kernel(const ScanParams p, int n_step, int tile_rows) {
    extern __shared__ __align__(16) float smem_scan[];
    const int T = p.T;
    const int t0 = blockIdx.y * tile_rows;
    const int rows_out = min(tile_rows, T - t0);
    const int rows_in = min(T - t0, tile_rows + n_step - 1);
    const int R = tile_rows + n_step - 1;  // allocated rows per array
    float *sr = smem_scan, *sd = sr + R * kNStepCols, *sv = sd + R * kNStepCols;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int b0 = blockIdx.x * kNStepCols;
    for (int ch = threadIdx.x; ch < rows_in * 32; ch += kNStepWarps * 32) {
        const int row = ch >> 5, c4 = (ch & 31) * 4;
        const int remaining = p.B - (b0 + c4);
        const int bytes = remaining >= 4 ? 16 : (remaining > 0 ? remaining * 4 : 0);
        const long long col = remaining > 0 ? b0 + c4 : 0;
        const long long g = (long long)(t0 + row) * p.st + col;
        cp_async16(sr + row * kNStepCols + c4, p.r + g, bytes);
        cp_async16(sd + row * kNStepCols + c4, p.disc + g, bytes);
        cp_async16(sv + row * kNStepCols + c4, p.v + g, bytes);
    }
    const int b = b0 + lane * 4;
    float vl[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) vl[q] = (b + q) < p.B ? __ldg(p.v + (long long)(T - 1) * p.st + b + q) : 0.f;
    cp_async_wait_all();
    __syncthreads();
    if (b >= p.B) return;
    const float4 *sr4 = reinterpret_cast<const float4 *>(sr), *sd4 = reinterpret_cast<const float4 *>(sd);
    const float4 *sv4 = reinterpret_cast<const float4 *>(sv);
    for (int tl = w; tl < rows_out; tl += kNStepWarps) {  // interleaved over warps
        const int t = t0 + tl, last = t + n_step - 1;
        float y[4];
        if (last < T) {
            const float4 v4 = sv4[(last - t0) * 32 + lane];
            y[0] = v4.x; y[1] = v4.y; y[2] = v4.z; y[3] = v4.w;
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) y[q] = vl[q];
        }
        for (int i = n_step - 1; i >= 0; --i) {
            const int kk = t + i;
            if (kk < T) {
                const int o = (kk - t0) * 32 + lane;
                const float4 r4 = sr4[o], d4 = sd4[o], v4 = sv4[o];
                y[0] = scan_step<MODE_LAMBDA_RETURNS>(r4.x, d4.x, v4.x, 0.f, p.lambda_, p.one_minus_lambda, y[0]);
                y[1] = scan_step<MODE_LAMBDA_RETURNS>(r4.y, d4.y, v4.y, 0.f, p.lambda_, p.one_minus_lambda, y[1]);
                y[2] = scan_step<MODE_LAMBDA_RETURNS>(r4.z, d4.z, v4.z, 0.f, p.lambda_, p.one_minus_lambda, y[2]);
                y[3] = scan_step<MODE_LAMBDA_RETURNS>(r4.w, d4.w, v4.w, 0.f, p.lambda_, p.one_minus_lambda, y[3]);
            } else {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    y[q] = scan_step<MODE_LAMBDA_RETURNS>(0.f, 1.f, vl[q], 0.f, p.lambda_, p.one_minus_lambda, y[q]);
            }
        }
        float *o = p.out + (long long)t * p.st + b;
        if (b + 3 < p.B) {
            __stcs(reinterpret_cast<float4 *>(o), make_float4(y[0], y[1], y[2], y[3]));
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (b + q < p.B) __stcs(o + q, y[q]);
        }
    }
}

static int32_t check_scan_layout(const char *who, int64_t B, int64_t T, int64_t sb, int64_t st) {
    B200RL_REQUIRE(B >= 1 && T >= 1, -2, "%s: need B>=1 and T>=1 (got B=%lld T=%lld)", who, (long long)B, (long long)T);
    B200RL_REQUIRE(B * (T + 1) < (1ll << 31), -3, "%s: B*T too large", who);
    B200RL_REQUIRE((st == 1 && sb >= T) || (sb == 1 && st >= B), -4,
                   "%s: strides (%lld,%lld) are neither batch-major (>=T,1) nor time-major (1,>=B)", who,
                   (long long)sb, (long long)st);
    return 0;
}

}  // namespace b200rl

extern "C" int32_t b200rl_lambda_returns(const float *r_t, const float *discount_t, const float *v_t,
                                         int64_t B, int64_t T, int64_t stride_b, int64_t stride_t,
                                         double lambda_, float *out, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(r_t && discount_t && v_t && out, -1, "lambda_returns: null pointer argument");
    if (int32_t rc = check_scan_layout("lambda_returns", B, T, stride_b, stride_t)) return rc;
    ScanParams p{r_t, discount_t, v_t, out, (int)B, (int)T, stride_b, stride_t, stride_b, stride_t, (float)lambda_,
                 (float)(1.0 - lambda_)};
    return launch_scan<MODE_LAMBDA_RETURNS>(p, (cudaStream_t)stream);
}

extern "C" int32_t b200rl_gae(const float *r_t, const float *discount_t, const float *values, int64_t B,
                              int64_t T, int64_t stride_b, int64_t stride_t, int64_t vstride_b,
                              int64_t vstride_t, double lambda_, float *out, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(r_t && discount_t && values && out, -1, "gae: null pointer argument");
    if (int32_t rc = check_scan_layout("gae", B, T, stride_b, stride_t)) return rc;
    B200RL_REQUIRE((vstride_t == 1 && vstride_b >= T + 1) || (vstride_b == 1 && vstride_t >= B), -4,
                   "gae: value strides (%lld,%lld) unsupported", (long long)vstride_b, (long long)vstride_t);
    // with one step or one column a [B,T] view is both batch- and time-major (strides (1,1)): nothing to compare then
    B200RL_REQUIRE((stride_b == 1) == (vstride_b == 1) || T == 1 || B == 1, -4,
                   "gae: values and rewards must share the major order");
    ScanParams p{r_t, discount_t, values, out, (int)B, (int)T, stride_b, stride_t, vstride_b, vstride_t, (float)lambda_,
                 (float)(1.0 - lambda_)};
    return launch_scan<MODE_GAE>(p, (cudaStream_t)stream);
}

extern "C" int32_t b200rl_n_step_returns(const float *r_t, const float *discount_t, const float *v_t, int64_t B,
                                         int64_t T, int64_t stride_b, int64_t stride_t, int64_t n,
                                         double lambda_, float *out, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(r_t && discount_t && v_t && out, -1, "n_step_returns: null pointer argument");
    if (int32_t rc = check_scan_layout("n_step_returns", B, T, stride_b, stride_t)) return rc;
    B200RL_REQUIRE(n >= 1 && n < (1 << 20), -2, "n_step_returns: n=%lld", (long long)n);
    ScanParams p{r_t, discount_t, v_t, out, (int)B, (int)T, stride_b, stride_t, stride_b, stride_t, (float)lambda_,
                 (float)(1.0 - lambda_)};
    const bool time_major = (stride_b == 1 && stride_t >= B);
    const bool al16 = ((reinterpret_cast<uintptr_t>(r_t) | reinterpret_cast<uintptr_t>(discount_t) |
                        reinterpret_cast<uintptr_t>(v_t) | reinterpret_cast<uintptr_t>(out)) % 16 == 0) && stride_t % 4 == 0;
    if (time_major && al16 && n - 1 <= kNStepMaxHalo) {
        int rows = env_int("B200RL_N4_ROWS", kNStepRows);
        rows = rows < 4 ? 4 : (rows > 64 ? 64 : rows);
        const dim3 grid((unsigned)((B + kNStepCols - 1) / kNStepCols), (unsigned)((T + rows - 1) / rows));
        const size_t smem = (size_t)3 * (rows + n - 1) * kNStepCols * 4;
        static size_t smem_cache[kMaxDevices];
        if (int32_t rc = ensure_dynamic_smem(n_step_smem_kernel, smem, smem_cache, "n_step_returns")) return rc;
        n_step_smem_kernel<<<grid, kNStepWarps * 32, smem, (cudaStream_t)stream>>>(p, (int)n, rows);
        return check_launch("n_step_smem_kernel");
    }
    const long long total = B * T;
    n_step_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(p, (int)n);
    return check_launch("n_step_kernel");
}
