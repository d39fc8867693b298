// K9: the memory-bound glue of the IMPALA-CNN between its convolutions (BASELINE config 5).
// The convolutions are K10/K11 (conv3x3.cu; cuDNN for 32 -> 32 channels); everything else of reference
// models/common.py:8-83 (ResidualBlock, ImpalaCNNBlock,
// ImpalaCNNLarge) -- bias adds, ReLUs, residual adds, 3x3/2 max-pools, their backward passes and the bias gradients --
// is HBM-bound byte shuffling that eager PyTorch spreads over ~25 % of the learner step (max_pool_backward_nhwc alone
// 8.7 ms of 41 ms at B=256, T=20).  All tensors are fp32 NHWC (torch channels_last), C % 4 == 0, one float4 = 4
// channels of one pixel, consecutive threads = consecutive float4 of memory.
//
//   obs_u8_to_f32_nhwc   x = obs.float() / 255 (reference models/models.py:73; as torch's CUDA kernel computes it:
//                        times the fp32 reciprocal) + NCHW -> NHWC, one pass
//   maxpool3x3s2_fwd     p = max_pool2d(u + bias[c], 3, 2, 1) (the bias of the preceding conv, added per tap so that ties
//                        break as in torch), optional second output relu(p), 1-byte argmax code (kh*3+kw) per element
//   maxpool3x3s2_bwd     gather form (a thread owns a 2x2 input patch and the 4 windows that cover it): no atomics,
//                        deterministic
//   bias_act_fwd         y = u + bias[c] (+ res);  optional outputs y and relu(y)
//   relu_bwd_add         g = dy + dr * (r > 0)  (either term optional), per-CTA per-channel sums of g -> partials
//                        (bias gradient of the op that produced the tensor; summed over CTAs by the caller in a fixed
//                        order -> deterministic)
#include "common.cuh"

namespace b200rl {

constexpr int kGlueThreads = 256;

__global__ void __launch_bounds__(kGlueThreads) obs_u8_to_f32_nhwc3_kernel(const uint8_t *__restrict__ obs,
                                                                           float *__restrict__ out, long long n_quads,
                                                                           long long quads_per_image, float divisor) {
    // one thread = 4 consecutive pixels of one image: 3 x uchar4 in (one per colour plane), 3 x float4 out
    const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_quads) return;
    const long long img = q / quads_per_image, qi = q - img * quads_per_image;
    const long long plane = quads_per_image * 4;
    const uchar4 *src = reinterpret_cast<const uchar4 *>(obs + img * 3 * plane) + qi;
    const uchar4 a = __ldg(src), b = __ldg(src + quads_per_image), c = __ldg(src + 2 * quads_per_image);
    float4 *dst = reinterpret_cast<float4 *>(out + (img * plane + qi * 4) * 3);
    // torch's CUDA `tensor / python_scalar` multiplies by the fp32 reciprocal (div_true_kernel_cuda); so do we
    const float d = 1.0f / divisor;
    __stcs(dst + 0, make_float4(__fmul_rn((float)a.x, d), __fmul_rn((float)b.x, d), __fmul_rn((float)c.x, d),
                                __fmul_rn((float)a.y, d)));
    __stcs(dst + 1, make_float4(__fmul_rn((float)b.y, d), __fmul_rn((float)c.y, d), __fmul_rn((float)a.z, d),
                                __fmul_rn((float)b.z, d)));
    __stcs(dst + 2, make_float4(__fmul_rn((float)c.z, d), __fmul_rn((float)a.w, d), __fmul_rn((float)b.w, d),
                                __fmul_rn((float)c.w, d)));
}

// frames whose H*W is not a multiple of 4 (or unaligned buffers): one thread per pixel, same arithmetic
__global__ void __launch_bounds__(kGlueThreads) obs_u8_to_f32_nhwc3_pixel_kernel(const uint8_t *__restrict__ obs,
                                                                                 float *__restrict__ out, long long n_pixels,
                                                                                 long long plane, float divisor) {
    const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_pixels) return;
    const long long img = q / plane, pi = q - img * plane;
    const uint8_t *src = obs + img * 3 * plane + pi;
    const float d = 1.0f / divisor;
    float *dst = out + q * 3;
    dst[0] = __fmul_rn((float)__ldg(src), d);
    dst[1] = __fmul_rn((float)__ldg(src + plane), d);
    dst[2] = __fmul_rn((float)__ldg(src + 2 * plane), d);
}

// torch's rule (max_pool_forward_nhwc): scan kh, kw ascending, take v if (v > best) || isnan(v) -> first maximum wins
__device__ __forceinline__ void pool_take(float v, int pos, float &best, uint32_t &arg) {
    if (v > best || v != v) { best = v; arg = (uint32_t)pos; }
}

// Pool kernels: block = (columns of one image row: ow * C4 + c4) x (rows), grid = (column blocks, row blocks, images).
// With C/4 a power of two (LOG_C4 >= 0) no thread divides by a run-time value; the first version flattened everything
// into one 64-bit index and spent 60-80 % of its issue slots on the divisions (ncu: profiles/r01_ncu_cnn_kernels_*).
template <int LOG_C4>
__device__ __forceinline__ void split_col(unsigned col, int C4, int &px, int &c4) {
    if constexpr (LOG_C4 >= 0) { px = (int)(col >> LOG_C4); c4 = (int)(col & ((1u << LOG_C4) - 1u)); }
    else { px = (int)(col / (unsigned)C4); c4 = (int)(col - (unsigned)px * (unsigned)C4); }
}

template <bool WITH_RELU, int LOG_C4>
__global__ void __launch_bounds__(kGlueThreads) maxpool3x3s2_fwd_kernel(
    const float4 *__restrict__ u, const float4 *__restrict__ bias, float4 *__restrict__ p, float4 *__restrict__ r,
    uint32_t *__restrict__ code, int N, int H, int W, int OH, int OW, int C4) {
    const unsigned col = blockIdx.x * blockDim.x + threadIdx.x;
    const int oh = blockIdx.y * blockDim.y + threadIdx.y;
    if (col >= (unsigned)(OW * C4) || oh >= OH) return;
    int ow, c4;
    split_col<LOG_C4>(col, C4, ow, c4);
    const float4 b = bias ? __ldg(bias + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
    for (int n = blockIdx.z; n < N; n += gridDim.z) {
        const float4 *base = u + (long long)n * H * W * C4 + c4;
        float4 best = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
        // the 9 taps are independent loads; windows are clipped at the border (padding 1: at least 4 taps in range)
        bool first = true;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int h = 2 * oh - 1 + kh;
            if (h < 0 || h >= H) continue;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const int w = 2 * ow - 1 + kw;
                if (w < 0 || w >= W) continue;
                float4 v = __ldg(base + (h * W + w) * C4);
                v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;  // per tap, so ties break exactly as max_pool2d(u + bias)
                const int pos = kh * 3 + kw;
                if (first) {  // torch initialises (best, arg) with the first in-range tap
                    best = v;
                    a0 = a1 = a2 = a3 = (uint32_t)pos;
                    first = false;
                } else {
                    pool_take(v.x, pos, best.x, a0);
                    pool_take(v.y, pos, best.y, a1);
                    pool_take(v.z, pos, best.z, a2);
                    pool_take(v.w, pos, best.w, a3);
                }
            }
        }
        const long long i = ((long long)n * OH + oh) * (OW * C4) + col;
        p[i] = best;
        if (WITH_RELU) r[i] = make_float4(fmaxf(best.x, 0.f), fmaxf(best.y, 0.f), fmaxf(best.z, 0.f), fmaxf(best.w, 0.f));
        code[i] = a0 | (a1 << 8) | (a2 << 16) | (a3 << 24);
    }
}

// Backward: one thread owns the 2x2 input patch (2k..2k+1, 2l..2l+1) of one channel quad.  The patch is covered by
// exactly the windows (k,l), (k,l+1), (k+1,l), (k+1,l+1), each fetched once (gradient + 4 tap codes); which tap of a
// window a patch pixel is, is a compile-time constant:
//   pixel (0,0): window (k,l) tap 4            pixel (0,1): (k,l) tap 5, (k,l+1) tap 3
//   pixel (1,0): (k,l) tap 7, (k+1,l) tap 1    pixel (1,1): (k,l) tap 8, (k,l+1) tap 6, (k+1,l) tap 2, (k+1,l+1) tap 0
// (the first version, one thread per input pixel, fetched 9 windows per patch and was issue-bound: ncu 75 %).
__device__ __forceinline__ void pool_acc(float4 &acc, const float4 &gv, uint32_t cw, uint32_t tap) {
    const uint32_t hit = __vcmpeq4(cw, tap * 0x01010101u);   // 0xff in every byte whose code is this tap
    if (hit & 0x000000ffu) acc.x += gv.x;
    if (hit & 0x0000ff00u) acc.y += gv.y;
    if (hit & 0x00ff0000u) acc.z += gv.z;
    if (hit & 0xff000000u) acc.w += gv.w;
}

template <int LOG_C4>
__global__ void __launch_bounds__(kGlueThreads) maxpool3x3s2_bwd_kernel(
    const float4 *__restrict__ g, const uint32_t *__restrict__ code, float4 *__restrict__ du, int N, int H, int W, int OH,
    int OW, int C4) {
    const unsigned col = blockIdx.x * blockDim.x + threadIdx.x;
    const int k = blockIdx.y * blockDim.y + threadIdx.y;
    if (col >= (unsigned)(OW * C4) || k >= OH) return;
    int l, c4;
    split_col<LOG_C4>(col, C4, l, c4);
    const bool down = k + 1 < OH, right = l + 1 < OW;          // windows (k+1, .), (., l+1) exist
    const bool row1 = 2 * k + 1 < H, col1 = 2 * l + 1 < W;     // patch pixels (1, .), (., 1) exist
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int n = blockIdx.z; n < N; n += gridDim.z) {
        const long long o00 = ((long long)n * OH + k) * (OW * C4) + col;
        const float4 g00 = __ldg(g + o00);
        const uint32_t c00 = __ldg(code + o00);
        float4 g01 = zero, g10 = zero, g11 = zero;
        uint32_t c01 = 0xffffffffu, c10 = 0xffffffffu, c11 = 0xffffffffu;   // 0xff matches no tap
        if (right) { g01 = __ldg(g + o00 + C4); c01 = __ldg(code + o00 + C4); }
        if (down) { g10 = __ldg(g + o00 + OW * C4); c10 = __ldg(code + o00 + OW * C4); }
        if (right && down) { g11 = __ldg(g + o00 + OW * C4 + C4); c11 = __ldg(code + o00 + OW * C4 + C4); }
        float4 *out = du + (((long long)n * H + 2 * k) * W + 2 * l) * C4 + c4;
        float4 acc = zero;
        pool_acc(acc, g00, c00, 4);
        __stcs(out, acc);
        if (col1) {
            acc = zero;
            pool_acc(acc, g00, c00, 5);
            pool_acc(acc, g01, c01, 3);
            __stcs(out + C4, acc);
        }
        if (row1) {
            acc = zero;
            pool_acc(acc, g00, c00, 7);
            pool_acc(acc, g10, c10, 1);
            __stcs(out + W * C4, acc);
            if (col1) {
                acc = zero;
                pool_acc(acc, g00, c00, 8);
                pool_acc(acc, g01, c01, 6);
                pool_acc(acc, g10, c10, 2);
                pool_acc(acc, g11, c11, 0);
                __stcs(out + W * C4 + C4, acc);
            }
        }
    }
}

// block / grid shape of the pool kernels: as many whole rows per 256-thread block as fit
static inline void pool_launch_shape(int cols, int rows, long long N, dim3 &block, dim3 &grid) {
    int bx = cols >= kGlueThreads ? kGlueThreads : ((cols + 31) / 32) * 32;
    int by = kGlueThreads / bx;
    if (by > rows) by = rows;
    block = dim3((unsigned)bx, (unsigned)by, 1);
    grid = dim3((unsigned)((cols + bx - 1) / bx), (unsigned)((rows + by - 1) / by), (unsigned)(N > 65535 ? 65535 : N));
}

static inline int log2_c4(int C4) { return C4 == 1 ? 0 : C4 == 2 ? 1 : C4 == 4 ? 2 : C4 == 8 ? 3 : C4 == 16 ? 4 : -1; }

template <bool HAS_RES, bool WRITE_Y, bool WRITE_R>
__global__ void __launch_bounds__(kGlueThreads) bias_act_fwd_kernel(const float4 *__restrict__ u,
                                                                    const float4 *__restrict__ bias,
                                                                    const float4 *__restrict__ res,
                                                                    float4 *__restrict__ y, float4 *__restrict__ r,
                                                                    long long total, int C4) {
    const long long stride = (long long)gridDim.x * blockDim.x;  // a multiple of C4: the channel quad of a thread is fixed
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const float4 b = __ldg(bias + (int)(i % C4));
#pragma unroll 4
    for (; i < total; i += stride) {
        float4 v = __ldcs(u + i);
        v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
        if (HAS_RES) {
            const float4 s = __ldcs(res + i);
            // torch order: x + (conv + bias)
            v.x = s.x + v.x; v.y = s.y + v.y; v.z = s.z + v.z; v.w = s.w + v.w;
        }
        if (WRITE_Y) y[i] = v;
        if (WRITE_R) r[i] = make_float4(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f), fmaxf(v.z, 0.f), fmaxf(v.w, 0.f));
    }
}

template <bool HAS_DY, bool HAS_DR, bool WRITE_G>
__global__ void __launch_bounds__(kGlueThreads) relu_bwd_add_kernel(const float4 *__restrict__ dy,
                                                                    const float4 *__restrict__ dr,
                                                                    const float4 *__restrict__ r,
                                                                    float4 *__restrict__ g, float *__restrict__ partials,
                                                                    long long total, int C4) {
    __shared__ float4 red[kGlueThreads];
    const long long stride = (long long)gridDim.x * blockDim.x;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (HAS_DR) {
            const float4 d = __ldcs(dr + i), m = __ldcs(r + i);
            v.x = m.x > 0.f ? d.x : 0.f; v.y = m.y > 0.f ? d.y : 0.f;
            v.z = m.z > 0.f ? d.z : 0.f; v.w = m.w > 0.f ? d.w : 0.f;
        }
        if (HAS_DY) {
            const float4 d = __ldcs(dy + i);
            v.x = d.x + v.x; v.y = d.y + v.y; v.z = d.z + v.z; v.w = d.w + v.w;
        }
        if (WRITE_G) g[i] = v;
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    // threads tid, tid + C4, tid + 2 C4, ... own the same channel quad: fixed-order sum by the first C4 threads
    red[threadIdx.x] = acc;
    __syncthreads();
    if ((int)threadIdx.x < C4) {
        float4 s = red[threadIdx.x];
        for (int k = threadIdx.x + C4; k < kGlueThreads; k += C4) {
            const float4 o = red[k];
            s.x += o.x; s.y += o.y; s.z += o.z; s.w += o.w;
        }
        reinterpret_cast<float4 *>(partials)[(long long)blockIdx.x * C4 + threadIdx.x] = s;
    }
}

static inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace b200rl

extern "C" int32_t b200rl_obs_u8_to_f32_nhwc(const uint8_t *obs, float *out, int64_t N, int64_t C, int64_t H, int64_t W,
                                              float divisor, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(N >= 0 && H >= 1 && W >= 1, -2, "obs_u8_to_f32_nhwc: N=%lld H=%lld W=%lld", (long long)N,
                   (long long)H, (long long)W);
    if (N == 0) return 0;
    B200RL_REQUIRE(obs && out, -1, "obs_u8_to_f32_nhwc: null pointer argument");
    B200RL_REQUIRE(C == 3, -2, "obs_u8_to_f32_nhwc: C=%lld (3 colour planes only)", (long long)C);
    B200RL_REQUIRE(divisor != 0.f, -2, "obs_u8_to_f32_nhwc: divisor 0");
    if ((H * W) % 4 != 0 || (reinterpret_cast<uintptr_t>(obs) & 3u) != 0 || !aligned16(out)) {
        const long long plane = H * W, np = N * plane;
        obs_u8_to_f32_nhwc3_pixel_kernel<<<(unsigned)((np + kGlueThreads - 1) / kGlueThreads), kGlueThreads, 0,
                                           (cudaStream_t)stream>>>(obs, out, np, plane, divisor);
        return check_launch("obs_u8_to_f32_nhwc3_pixel_kernel");
    }
    const long long qpi = H * W / 4, nq = N * qpi;
    obs_u8_to_f32_nhwc3_kernel<<<(unsigned)((nq + kGlueThreads - 1) / kGlueThreads), kGlueThreads, 0,
                                 (cudaStream_t)stream>>>(obs, out, nq, qpi, divisor);
    return check_launch("obs_u8_to_f32_nhwc3_kernel");
}

extern "C" int32_t b200rl_maxpool3x3s2_fwd(const float *u, const float *bias, float *p, float *r, uint8_t *code,
                                            int64_t N, int64_t H, int64_t W, int64_t C, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(N >= 0 && N < (1ll << 31) && H >= 1 && W >= 1 && H < (1 << 20) && W < (1 << 20), -2, "maxpool3x3s2_fwd: N=%lld H=%lld W=%lld",
                   (long long)N, (long long)H, (long long)W);
    B200RL_REQUIRE(C >= 4 && C % 4 == 0 && C <= 4096, -2, "maxpool3x3s2_fwd: C=%lld must be a multiple of 4", (long long)C);
    if (N == 0) return 0;
    B200RL_REQUIRE(u && p && code, -1, "maxpool3x3s2_fwd: null pointer argument");
    B200RL_REQUIRE(aligned16(u) && aligned16(p) && aligned16(r) && aligned16(bias) && (reinterpret_cast<uintptr_t>(code) & 3u) == 0,
                   -3, "maxpool3x3s2_fwd: tensors must be 16-byte aligned");
    const int OH = (int)((H - 1) / 2 + 1), OW = (int)((W - 1) / 2 + 1), C4 = (int)(C / 4);
    B200RL_REQUIRE(H * W * C4 < (1ll << 31), -2, "maxpool3x3s2_fwd: image of %lld float4 is too large", (long long)(H * W * C4));
    dim3 block, grid;
    pool_launch_shape(OW * C4, OH, N, block, grid);
    cudaStream_t s = (cudaStream_t)stream;
#define B200RL_PF(R, L) maxpool3x3s2_fwd_kernel<R, L><<<grid, block, 0, s>>>( \
        (const float4 *)u, (const float4 *)bias, (float4 *)p, (float4 *)r, (uint32_t *)code, (int)N, (int)H, (int)W, OH, OW, C4)
#define B200RL_PFL(R) switch (log2_c4(C4)) { case 0: B200RL_PF(R, 0); break; case 1: B200RL_PF(R, 1); break; \
        case 2: B200RL_PF(R, 2); break; case 3: B200RL_PF(R, 3); break; case 4: B200RL_PF(R, 4); break; default: B200RL_PF(R, -1); }
    if (r) { B200RL_PFL(true) } else { B200RL_PFL(false) }
#undef B200RL_PFL
#undef B200RL_PF
    return check_launch("maxpool3x3s2_fwd_kernel");
}

extern "C" int32_t b200rl_maxpool3x3s2_bwd(const float *g, const uint8_t *code, float *du, int64_t N, int64_t H,
                                            int64_t W, int64_t C, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(N >= 0 && N < (1ll << 31) && H >= 1 && W >= 1 && H < (1 << 20) && W < (1 << 20), -2, "maxpool3x3s2_bwd: N=%lld H=%lld W=%lld",
                   (long long)N, (long long)H, (long long)W);
    B200RL_REQUIRE(C >= 4 && C % 4 == 0 && C <= 4096, -2, "maxpool3x3s2_bwd: C=%lld must be a multiple of 4", (long long)C);
    if (N == 0) return 0;
    B200RL_REQUIRE(g && code && du, -1, "maxpool3x3s2_bwd: null pointer argument");
    B200RL_REQUIRE(aligned16(g) && aligned16(du) && (reinterpret_cast<uintptr_t>(code) & 3u) == 0, -3,
                   "maxpool3x3s2_bwd: tensors must be 16-byte aligned");
    const int OH = (int)((H - 1) / 2 + 1), OW = (int)((W - 1) / 2 + 1), C4 = (int)(C / 4);
    B200RL_REQUIRE(H * W * C4 < (1ll << 31), -2, "maxpool3x3s2_bwd: image of %lld float4 is too large", (long long)(H * W * C4));
    dim3 block, grid;
    pool_launch_shape(OW * C4, OH, N, block, grid);            // one thread per 2x2 input patch and channel quad
    cudaStream_t s = (cudaStream_t)stream;
#define B200RL_PB(L) maxpool3x3s2_bwd_kernel<L><<<grid, block, 0, s>>>( \
        (const float4 *)g, (const uint32_t *)code, (float4 *)du, (int)N, (int)H, (int)W, OH, OW, C4)
    switch (log2_c4(C4)) { case 0: B200RL_PB(0); break; case 1: B200RL_PB(1); break; case 2: B200RL_PB(2); break;
        case 3: B200RL_PB(3); break; case 4: B200RL_PB(4); break; default: B200RL_PB(-1); }
#undef B200RL_PB
    return check_launch("maxpool3x3s2_bwd_kernel");
}

namespace b200rl {
static inline int32_t glue_check(const char *what, int64_t n_pix, int64_t C, int64_t n_ctas) {
    B200RL_REQUIRE(n_pix >= 0, -2, "%s: n_pix=%lld", what, (long long)n_pix);
    B200RL_REQUIRE(C >= 4 && C % 4 == 0 && kGlueThreads % (C / 4) == 0, -2,
                   "%s: C=%lld (C/4 must divide %d)", what, (long long)C, kGlueThreads);
    B200RL_REQUIRE(n_ctas >= 1 && n_ctas <= 65535, -2, "%s: n_ctas=%lld", what, (long long)n_ctas);
    return 0;
}
}  // namespace b200rl

extern "C" int32_t b200rl_bias_act_fwd(const float *u, const float *bias, const float *res, float *y, float *r,
                                        int64_t n_pix, int64_t C, int64_t n_ctas, void *stream) {
    using namespace b200rl;
    if (int32_t e = glue_check("bias_act_fwd", n_pix, C, n_ctas)) return e;
    if (n_pix == 0) return 0;
    B200RL_REQUIRE(u && bias && (y || r), -1, "bias_act_fwd: null pointer argument");
    B200RL_REQUIRE(aligned16(u) && aligned16(bias) && aligned16(res) && aligned16(y) && aligned16(r), -3,
                   "bias_act_fwd: tensors must be 16-byte aligned");
    const int C4 = (int)(C / 4);
    const long long total = n_pix * C4;
    const unsigned grid = (unsigned)n_ctas;
    cudaStream_t s = (cudaStream_t)stream;
#define B200RL_BA(R, Y, Q) bias_act_fwd_kernel<R, Y, Q><<<grid, kGlueThreads, 0, s>>>( \
        (const float4 *)u, (const float4 *)bias, (const float4 *)res, (float4 *)y, (float4 *)r, total, C4)
    if (res) {
        if (y && r) B200RL_BA(true, true, true); else if (y) B200RL_BA(true, true, false); else B200RL_BA(true, false, true);
    } else {
        if (y && r) B200RL_BA(false, true, true); else if (y) B200RL_BA(false, true, false); else B200RL_BA(false, false, true);
    }
#undef B200RL_BA
    return check_launch("bias_act_fwd_kernel");
}

extern "C" int32_t b200rl_relu_bwd_add(const float *dy, const float *dr, const float *r, float *g, float *partials,
                                        int64_t n_pix, int64_t C, int64_t n_ctas, void *stream) {
    using namespace b200rl;
    if (int32_t e = glue_check("relu_bwd_add", n_pix, C, n_ctas)) return e;
    B200RL_REQUIRE(partials, -1, "relu_bwd_add: null partials");
    B200RL_REQUIRE((dy || dr) && (!dr || r), -1, "relu_bwd_add: needs dy and/or (dr and r)");
    B200RL_REQUIRE(g || !dr, -1, "relu_bwd_add: g is required when dr is given");
    B200RL_REQUIRE(aligned16(dy) && aligned16(dr) && aligned16(r) && aligned16(g) && aligned16(partials), -3,
                   "relu_bwd_add: tensors must be 16-byte aligned");
    const int C4 = (int)(C / 4);
    const long long total = n_pix * C4;
    const unsigned grid = (unsigned)n_ctas;
    cudaStream_t s = (cudaStream_t)stream;
#define B200RL_RB(Y, R, G) relu_bwd_add_kernel<Y, R, G><<<grid, kGlueThreads, 0, s>>>( \
        (const float4 *)dy, (const float4 *)dr, (const float4 *)r, (float4 *)g, partials, total, C4)
    if (dy && dr) B200RL_RB(true, true, true);
    else if (dr) B200RL_RB(false, true, true);
    else if (g) B200RL_RB(true, false, true);
    else B200RL_RB(true, false, false);  // column sums only
#undef B200RL_RB
    return check_launch("relu_bwd_add_kernel");
}
