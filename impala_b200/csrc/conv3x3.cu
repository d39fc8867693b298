// K10: 3x3 / stride 1 / padding 1 convolution weight gradient for the IMPALA-CNN's small channel counts
// (reference models/common.py:12-26,41-45: every convolution of ImpalaCNNLarge is 3x3, pad 1, with 3/16/32 input and
// 16/32 output channels).  fp32 NHWC activations, TF32 tensor-core products with fp32 accumulation -- the arithmetic
// torch's default (cudnn.allow_tf32 = True) already uses for these layers.
//
//   dW[co][kh][kw][ci] = sum over (n, h, w) of dy[n,h,w,co] * x[n, h+kh-1, w+kw-1, ci]
//
// is a GEMM with a tiny output (<= 32 x 288) and a reduction over millions of pixels: each pixel's Cin + Cout floats
// feed Cout*9*Cin MACs, so the layer is HBM-bound once the products run on tensor cores (7.3 GB over the 15 layers
// at B=256, T=20: ~1.1 ms; cuDNN's wgrad kernels take 10.1 ms).  Layout of the work:
//   * one persistent CTA per SM walks tiles of 8 image rows (tile = blockIdx, += gridDim); the x rows with a one-pixel
//     halo and the dy rows are staged in shared memory by cp.async in a 3-buffer ring: the copies of tiles k+1 and
//     k+2 fly while tile k is multiplied, one __syncthreads per tile;
//   * 24 (32 -> 32 channels: 12) warps = 3 kernel rows (kh) x 8 (4) pixel groups; a warp takes octets of 8 consecutive
//     pixels of a row as the K dimension of mma.sync.m16n8k8 (A = dy^T: Cout x 8 pixels, B = shifted x: 8 pixels x
//     Cin) for kw = 0..2, and keeps its 3 x (Cout/16) x (Cin/8) accumulator tiles in registers across ALL tiles of
//     the CTA; the image width is a template parameter, so every fragment address is base + immediate (the first
//     version spent 10 integer instructions per MMA on address arithmetic: ncu, 67 % issue-slot utilisation);
//   * pixel strides in shared memory are = 8 (mod 32) floats, which makes every fragment load conflict-free;
//   * at the end the pixel groups are added through shared memory in a fixed order and the CTA writes ONE partial
//     dW; the caller adds the per-CTA partials in order (deterministic, no atomics).
#include "common.cuh"

namespace b200rl {

constexpr int kWgTH = 8;            // image rows per tile
// warps = 3 kernel rows x pixel groups; the 32 -> 32 layers keep 96 accumulator registers per thread, so 12 warps
template <int CIN_P, int COUT>
struct WgWarps { static constexpr int value = (CIN_P == 32 && COUT == 32) ? 12 : 24; };

__device__ __forceinline__ void cp_async4(void *smem, const void *gmem, int src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(smem_u32(smem)), "l"(gmem), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// TF32 operands: fp32 bit patterns (the tensor core reads the sign, the exponent and the top 10 mantissa bits), rounded
// to nearest on the way from shared memory into the fragment registers (tf32_rna below: one integer add).
__device__ __forceinline__ void mma_tf32_16x8x8(float (&d)[4], const float (&a)[4], float b0, float b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(__float_as_uint(a[0])), "r"(__float_as_uint(a[1])), "r"(__float_as_uint(a[2])), "r"(__float_as_uint(a[3])),
          "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)));
}

template <int CIN_P>
struct WgStride { static constexpr int value = CIN_P == 8 ? 8 : CIN_P + 8; };   // = 8 (mod 32)

// Asynchronous staging of one tile: x rows h0-1 .. h0+TH (zero outside the image) and dy rows h0 .. h0+TH-1.
// A thread owns one 16-byte (Cin = 3: 4-byte) column chunk of the row and walks the rows, so the per-copy cost is an
// add and a bounds test; the chunk's offsets are computed once per kernel (RowPlan).  Rows whose chunk count does
// not divide the CTA size take the generic (divide per copy) loop.
struct RowPlan {
    int soff, goff, r0, rstep;   // shared / global offset of the chunk inside a row (floats), first row, row step
    bool fast, active;
};

template <int THREADS>
__device__ __forceinline__ RowPlan make_plan(int per_row, int floats_per_chunk, int C, int smem_pixel_stride, int col0) {
    RowPlan p;
    p.fast = (THREADS % per_row) == 0;
    const int c = threadIdx.x % per_row, e = c * floats_per_chunk, px = e / C;
    p.soff = (px + col0) * smem_pixel_stride + (e - px * C);
    p.goff = e;
    p.r0 = threadIdx.x / per_row;
    p.rstep = THREADS / per_row;
    p.active = true;
    return p;
}

__device__ __forceinline__ void cp_async_u32(uint32_t saddr, const void *gmem, int bytes16, int src_bytes) {
    if (bytes16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(saddr), "l"(gmem), "r"(src_bytes) : "memory");
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(saddr), "l"(gmem), "r"(src_bytes) : "memory");
}

template <int BYTES, int THREADS>
__device__ __forceinline__ void stage_rows(float *s, int s_row_stride, const float *__restrict__ g, long long g_row_stride,
                                           long long image_row0, int h_first, int rows, int H, const RowPlan &p,
                                           int per_row, int C, int smem_pixel_stride, int col0) {
    // image_row0 = n * H; row r of the tile is image row h_first + r
    if (p.fast) {
        // one pointer / one shared address per thread, stepped by whole rows: ~6 instructions per copy
        const float *src = g + (image_row0 + h_first + p.r0) * g_row_stride + p.goff;   // dereferenced only when in range
        uint32_t dst = smem_u32(s) + 4u * (unsigned)(p.r0 * s_row_stride + p.soff);
        const long long gstep = (long long)p.rstep * g_row_stride;
        const uint32_t sstep = 4u * (unsigned)(p.rstep * s_row_stride);
        for (int r = p.r0; r < rows; r += p.rstep, src += gstep, dst += sstep) {
            const bool in = (unsigned)(h_first + r) < (unsigned)H;
            cp_async_u32(dst, in ? src : g, BYTES == 16, in ? BYTES : 0);
        }
    } else {
        constexpr int FPC = BYTES / 4;
        for (int i = threadIdx.x; i < rows * per_row; i += THREADS) {
            const int r = i / per_row, e = (i - r * per_row) * FPC, px = e / C, h = h_first + r;
            const bool in = h >= 0 && h < H;
            const float *src = g + (image_row0 + (in ? h : 0)) * g_row_stride + e;
            float *dst = s + r * s_row_stride + (px + col0) * smem_pixel_stride + (e - px * C);
            cp_async_u32(smem_u32(dst), src, BYTES == 16, in ? BYTES : 0);
        }
    }
}

// fp32 -> TF32 with round-to-nearest, ties away from zero (what cvt.rna.tf32.f32 computes): the tensor core drops the
// low 13 mantissa bits, so adding half an ulp of the kept precision to the bit pattern beforehand rounds.  Feeding
// the raw bits (truncation) shrinks every product by ~2^-11 on average -- coherently, so a 15-layer network drifted
// by 1.6 % against cuDNN's 0.3 % (tests/test_gpu_cnn_glue.py::test_fused_model_matches_eager_model).
__device__ __forceinline__ float tf32_rna(float f) { return __uint_as_float(__float_as_uint(f) + 0x1000u); }

// ring depth: 3 buffers where they fit in 227 KB, else 2
template <int CIN_P, int COUT, int W>
struct WgStages {
    static constexpr int tile_bytes = 4 * ((kWgTH + 2) * (W + 2) * WgStride<CIN_P>::value + kWgTH * W * (COUT + 8));
    static constexpr int value = 3 * tile_bytes <= 227 * 1024 ? 3 : 2;
};

template <int CIN_P, int COUT, int W>
__global__ void __launch_bounds__(WgWarps<CIN_P, COUT>::value * 32, 1)
conv3x3_wgrad_kernel(const float *__restrict__ x, const float *__restrict__ dy, float *__restrict__ partials, int N, int H,
                     int Cin, int n_tiles) {
    constexpr int XS = WgStride<CIN_P>::value, DS = COUT + 8;
    constexpr int MT = COUT / 16, NT = CIN_P / 8, WARPS = WgWarps<CIN_P, COUT>::value, THREADS = WARPS * 32, PG = WARPS / 3;
    constexpr int XBYTES = CIN_P == 8 ? 4 : 16;             // Cin = 3: 12-byte pixels, 4-byte copies
    constexpr int XW = W + 2;                               // tile columns incl. the halo
    constexpr int x_floats = (kWgTH + 2) * XW * XS, tile_floats = x_floats + kWgTH * W * DS;
    constexpr int OCTS_PER_ROW = W / 8, N_OCTS = kWgTH * OCTS_PER_ROW, kWgStages = WgStages<CIN_P, COUT, W>::value;
    static_assert(N_OCTS % PG == 0, "pixel groups must divide the octets of a tile");
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, gid = lane >> 2, tig = lane & 3;
    const int kh = warp % 3, pg = warp / 3;
    const int tiles_per_image = (H + kWgTH - 1) / kWgTH;
    const int x_per_row = W * Cin / (XBYTES / 4);
    constexpr int dy_per_row = W * COUT / 4;
    const RowPlan xplan = make_plan<THREADS>(x_per_row, XBYTES / 4, Cin, XS, 1);
    const RowPlan dplan = make_plan<THREADS>(dy_per_row, 4, COUT, DS, 0);

    float acc[3][MT][NT][4];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < MT; ++b)
#pragma unroll
            for (int c = 0; c < NT; ++c)
#pragma unroll
                for (int d = 0; d < 4; ++d) acc[a][b][c][d] = 0.f;

    // zero once, in every buffer, what the tile copies never touch: halo columns, padded input channels
    for (int i = tid; i < kWgStages * tile_floats; i += THREADS) smem[i] = 0.f;
    __syncthreads();

    auto stage = [&](int t, int b) {
        if (t < n_tiles) {
            float *xs = smem + b * tile_floats;
            const int n = t / tiles_per_image, h0 = (t - n * tiles_per_image) * kWgTH;
            stage_rows<XBYTES, THREADS>(xs, XW * XS, x, (long long)W * Cin, (long long)n * H, h0 - 1, kWgTH + 2, H, xplan,
                                        x_per_row, Cin, XS, 1);
            stage_rows<16, THREADS>(xs + x_floats, W * DS, dy, (long long)W * COUT, (long long)n * H, h0, kWgTH, H, dplan,
                                    dy_per_row, COUT, DS, 0);
        }
        cp_async_commit();                                  // one group per round, empty past the last tile
    };

    // per-warp fragment bases inside a tile (all remaining offsets are compile-time constants)
    const int a_base = x_floats + tig * DS + gid;           // dy: pixel tig, channel gid
    const int b_base = (kh * XW + tig) * XS + gid;          // x: tile row kh, column tig (= w - 1 + kw at kw = 0)

    int tile = blockIdx.x, buf = 0;
#pragma unroll
    for (int k = 0; k < kWgStages - 1; ++k) stage(tile + k * gridDim.x, k);
    for (; tile < n_tiles; tile += gridDim.x, buf = (buf + 1 == kWgStages ? 0 : buf + 1)) {
        cp_async_wait_group<kWgStages - 2>();               // everything but the newest group(s): this tile has landed
        __syncthreads();                                    // ... for every thread; and the buffer consumed in the
        stage(tile + (kWgStages - 1) * gridDim.x, buf == 0 ? kWgStages - 1 : buf - 1);   // previous round is free to refill
        const float *tb = smem + buf * tile_floats;
        // ---- products: this warp's kernel row, its share of the pixel octets, all output channels ----
#pragma unroll
        for (int k = 0; k < N_OCTS / PG; ++k) {
            const int oct = pg + k * PG, row = oct / OCTS_PER_ROW, w0 = (oct % OCTS_PER_ROW) * 8;
            const float *dp = tb + a_base + (row * W + w0) * DS;
            float a[MT][4];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) {
                a[mt][0] = tf32_rna(dp[mt * 16]);
                a[mt][1] = tf32_rna(dp[mt * 16 + 8]);
                a[mt][2] = tf32_rna(dp[4 * DS + mt * 16]);
                a[mt][3] = tf32_rna(dp[4 * DS + mt * 16 + 8]);
            }
            const float *xp = tb + b_base + (row * XW + w0) * XS;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const float b0 = tf32_rna(xp[kw * XS + nt * 8]), b1 = tf32_rna(xp[(kw + 4) * XS + nt * 8]);
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) mma_tf32_16x8x8(acc[kw][mt][nt], a[mt], b0, b1);
                }
            }
        }
    }
    cp_async_wait_group<0>();
    __syncthreads();

    // ---- add the pixel groups in order through shared memory: out[co][kh][kw][ci] (ci < CIN_P) ----
    float *out = smem;
    for (int g = 0; g < PG; ++g) {
        if (pg == g) {
#pragma unroll
            for (int kw = 0; kw < 3; ++kw)
#pragma unroll
                for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                        for (int d = 0; d < 4; ++d) {
                            const int co = mt * 16 + gid + ((d & 2) ? 8 : 0), ci = nt * 8 + 2 * tig + (d & 1);
                            float *o = out + ((co * 3 + kh) * 3 + kw) * CIN_P + ci;
                            *o = (g == 0 ? 0.f : *o) + acc[kw][mt][nt][d];
                        }
        }
        __syncthreads();
    }
    float *dst = partials + (long long)blockIdx.x * COUT * 9 * Cin;
    for (int i = tid; i < COUT * 9 * Cin; i += THREADS) {
        const int ci = i % Cin, rest = i / Cin;
        dst[i] = out[rest * CIN_P + ci];
    }
}

// ---- K10, row-streaming variant for the wide, few-channel layers (W >= 32, at most 16 input channels) ----------------
// In the kernel above a fragment of x feeds ONE MMA per output-channel tile, so at 16 channels the shared-memory pipe
// (2.7 fragment words per MMA) binds before the tensor pipe does.  Here a warp owns a strip of 8 columns x 4 dy rows
// and all three kernel rows of KWN column shifts: walking down the strip, the x fragment of tile row R meets the dy
// rows R, R-1, R-2 as taps kh = 0, 1, 2 -- three MMAs per x fragment, with the dy fragments of the last three rows kept
// in registers.  Two shapes, chosen by measurement (profiles/r01_conv_layers_bench.txt):
//   W = 32: a warp owns all 3 shifts (0.9 fragment words per MMA, 72-144 accumulator registers, 8 warps):
//           16 -> 16 channels 0.26 -> 0.21 ms, 16 -> 32 channels 0.47 -> 0.32 ms;
//   W = 64: one shift per warp (1.3 words per MMA, 24 warps): 3 -> 16 channels 0.60 -> 0.54 ms.
template <int CIN_P, int COUT, int W>
struct Wg2 {
    static constexpr int SPLIT = W == 64 ? 3 : 1;           // warps sharing a strip, each with 3 / SPLIT column shifts
    static constexpr int RS = W == 32 ? 4 : 8;              // dy rows per warp
    static constexpr bool use = W >= 32 && CIN_P <= 16 && (W == 64 || (COUT / 16) * (CIN_P / 8) <= 4);
    static constexpr int warps = (W / 8) * (kWgTH / RS) * SPLIT;
};

template <int CIN_P, int COUT, int W>
__global__ void __launch_bounds__(Wg2<CIN_P, COUT, W>::warps * 32, 1)
conv3x3_wgrad_rows_kernel(const float *__restrict__ x, const float *__restrict__ dy, float *__restrict__ partials, int N, int H,
                          int Cin, int n_tiles) {
    constexpr int XS = WgStride<CIN_P>::value, DS = COUT + 8;
    constexpr int MT = COUT / 16, NT = CIN_P / 8, WARPS = Wg2<CIN_P, COUT, W>::warps, THREADS = WARPS * 32;
    constexpr int XBYTES = CIN_P == 8 ? 4 : 16;
    constexpr int XW = W + 2;
    constexpr int x_floats = (kWgTH + 2) * XW * XS, tile_floats = x_floats + kWgTH * W * DS;
    constexpr int kWgStages = WgStages<CIN_P, COUT, W>::value, kWg2RS = Wg2<CIN_P, COUT, W>::RS;
    constexpr int SPLIT = Wg2<CIN_P, COUT, W>::SPLIT, KWN = 3 / SPLIT;
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, gid = lane >> 2, tig = lane & 3;
    const int kw0 = (warp % SPLIT) * KWN, strip = (warp / SPLIT) % (W / 8), seg = warp / (SPLIT * (W / 8));
    const int tiles_per_image = (H + kWgTH - 1) / kWgTH;
    const int x_per_row = W * Cin / (XBYTES / 4);
    constexpr int dy_per_row = W * COUT / 4;
    const RowPlan xplan = make_plan<THREADS>(x_per_row, XBYTES / 4, Cin, XS, 1);
    const RowPlan dplan = make_plan<THREADS>(dy_per_row, 4, COUT, DS, 0);

    float acc[3][KWN][MT][NT][4];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < KWN; ++b)
#pragma unroll
            for (int c = 0; c < MT; ++c)
#pragma unroll
                for (int d = 0; d < NT; ++d)
#pragma unroll
                    for (int e = 0; e < 4; ++e) acc[a][b][c][d][e] = 0.f;

    for (int i = tid; i < kWgStages * tile_floats; i += THREADS) smem[i] = 0.f;
    __syncthreads();

    auto stage = [&](int t, int b) {
        if (t < n_tiles) {
            float *xs = smem + b * tile_floats;
            const int n = t / tiles_per_image, h0 = (t - n * tiles_per_image) * kWgTH;
            stage_rows<XBYTES, THREADS>(xs, XW * XS, x, (long long)W * Cin, (long long)n * H, h0 - 1, kWgTH + 2, H, xplan,
                                        x_per_row, Cin, XS, 1);
            stage_rows<16, THREADS>(xs + x_floats, W * DS, dy, (long long)W * COUT, (long long)n * H, h0, kWgTH, H, dplan,
                                    dy_per_row, COUT, DS, 0);
        }
        cp_async_commit();
    };

    // fragment bases of this warp inside a tile: dy row seg*RS, column strip*8 + tig; x tile row seg*RS, column strip*8 + tig
    const int a_base = x_floats + ((seg * kWg2RS) * W + strip * 8 + tig) * DS + gid;
    const int b_base = ((seg * kWg2RS) * XW + strip * 8 + tig + kw0) * XS + gid;

    int tile = blockIdx.x, buf = 0;
#pragma unroll
    for (int k = 0; k < kWgStages - 1; ++k) stage(tile + k * gridDim.x, k);
    for (; tile < n_tiles; tile += gridDim.x, buf = (buf + 1 == kWgStages ? 0 : buf + 1)) {
        cp_async_wait_group<kWgStages - 2>();
        __syncthreads();
        stage(tile + (kWgStages - 1) * gridDim.x, buf == 0 ? kWgStages - 1 : buf - 1);
        const float *tb = smem + buf * tile_floats;
        float a[3][MT][4];                                   // dy fragments of the last three rows (slot = row % 3)
#pragma unroll
        for (int j = 0; j < kWg2RS + 2; ++j) {               // x tile row seg*RS + j  (image row h0 - 1 + seg*RS + j)
            float b[KWN][NT][2];
            const float *xp = tb + b_base + j * XW * XS;
#pragma unroll
            for (int k = 0; k < KWN; ++k)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    b[k][nt][0] = tf32_rna(xp[k * XS + nt * 8]);
                    b[k][nt][1] = tf32_rna(xp[(k + 4) * XS + nt * 8]);
                }
            if (j < kWg2RS) {
                const float *dp = tb + a_base + j * W * DS;
#pragma unroll
                for (int mt = 0; mt < MT; ++mt) {
                    a[j % 3][mt][0] = tf32_rna(dp[mt * 16]);
                    a[j % 3][mt][1] = tf32_rna(dp[mt * 16 + 8]);
                    a[j % 3][mt][2] = tf32_rna(dp[4 * DS + mt * 16]);
                    a[j % 3][mt][3] = tf32_rna(dp[4 * DS + mt * 16 + 8]);
                }
            }
#pragma unroll
            for (int kh = 0; kh < 3; ++kh) {
                const int jr = j - kh;                       // dy row (local) that sees x row j as kernel row kh
                if (jr < 0 || jr >= kWg2RS) continue;
#pragma unroll
                for (int k = 0; k < KWN; ++k)
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt)
                            mma_tf32_16x8x8(acc[kh][k][mt][nt], a[jr % 3][mt], b[k][nt][0], b[k][nt][1]);
            }
        }
    }
    cp_async_wait_group<0>();
    __syncthreads();

    // ---- add the warps in order through shared memory: out[co][kh][kw][ci] (ci < CIN_P) ----
    float *out = smem;
    for (int g = 0; g < WARPS; ++g) {
        if (warp == g) {
#pragma unroll
            for (int kh = 0; kh < 3; ++kh)
#pragma unroll
                for (int k = 0; k < KWN; ++k)
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                            for (int d = 0; d < 4; ++d) {
                                const int co = mt * 16 + gid + ((d & 2) ? 8 : 0), ci = nt * 8 + 2 * tig + (d & 1);
                                float *o = out + ((co * 3 + kh) * 3 + kw0 + k) * CIN_P + ci;
                                *o = (g < SPLIT ? 0.f : *o) + acc[kh][k][mt][nt][d];   // the first SPLIT warps open their shifts
                            }
        }
        __syncthreads();
    }
    float *dst = partials + (long long)blockIdx.x * COUT * 9 * Cin;
    for (int i = tid; i < COUT * 9 * Cin; i += THREADS) {
        const int ci = i % Cin, rest = i / Cin;
        dst[i] = out[rest * CIN_P + ci];
    }
}

template <int CIN_P, int COUT, int W>
static int32_t launch_wgrad(const float *x, const float *dy, float *partials, int N, int H, int Cin, int n_ctas,
                            cudaStream_t stream) {
    constexpr int XS = WgStride<CIN_P>::value, DS = COUT + 8;
    constexpr int tile_floats = (kWgTH + 2) * (W + 2) * XS + kWgTH * W * DS, out_floats = COUT * 9 * CIN_P;
    constexpr int kWgStages = WgStages<CIN_P, COUT, W>::value;
    constexpr int smem = 4 * (kWgStages * tile_floats > out_floats ? kWgStages * tile_floats : out_floats);
    if (smem > 227 * 1024) {
        set_error("conv3x3_wgrad: Cin=%d Cout=%d W=%d needs %d bytes of shared memory", Cin, COUT, W, smem);
        return -2;
    }
    static int configured[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    const int n_tiles = N * ((H + kWgTH - 1) / kWgTH);
    if constexpr (Wg2<CIN_P, COUT, W>::use) {
        if (dev < 64 && !configured[dev]) {
            if (cudaFuncSetAttribute(conv3x3_wgrad_rows_kernel<CIN_P, COUT, W>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     smem <= 227 * 1024 ? smem : 0) != cudaSuccess)
                return check_launch("conv3x3_wgrad: shared-memory opt-in");
            configured[dev] = 1;
        }
        conv3x3_wgrad_rows_kernel<CIN_P, COUT, W><<<n_ctas, Wg2<CIN_P, COUT, W>::warps * 32, smem, stream>>>(
            x, dy, partials, N, H, Cin, n_tiles);
        return check_launch("conv3x3_wgrad_rows_kernel");
    } else {
        if (dev < 64 && !configured[dev]) {
            if (cudaFuncSetAttribute(conv3x3_wgrad_kernel<CIN_P, COUT, W>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     smem <= 227 * 1024 ? smem : 0) != cudaSuccess)
                return check_launch("conv3x3_wgrad: shared-memory opt-in");
            configured[dev] = 1;
        }
        conv3x3_wgrad_kernel<CIN_P, COUT, W><<<n_ctas, WgWarps<CIN_P, COUT>::value * 32, smem, stream>>>(x, dy, partials, N,
                                                                                                      H, Cin, n_tiles);
        return check_launch("conv3x3_wgrad_kernel");
    }
}

template <int CIN_P, int COUT>
static int32_t launch_wgrad_w(const float *x, const float *dy, float *partials, int N, int H, int W, int Cin, int n_ctas,
                              cudaStream_t stream) {
    switch (W) {
        case 8: return launch_wgrad<CIN_P, COUT, 8>(x, dy, partials, N, H, Cin, n_ctas, stream);
        case 16: return launch_wgrad<CIN_P, COUT, 16>(x, dy, partials, N, H, Cin, n_ctas, stream);
        case 32: return launch_wgrad<CIN_P, COUT, 32>(x, dy, partials, N, H, Cin, n_ctas, stream);
        case 64: return launch_wgrad<CIN_P, COUT, 64>(x, dy, partials, N, H, Cin, n_ctas, stream);
    }
    set_error("conv3x3_wgrad: W=%d (supported: 8, 16, 32, 64)", W);
    return -2;
}

}  // namespace b200rl

extern "C" int32_t b200rl_conv3x3_wgrad(const float *x, const float *dy, float *partials, int64_t N, int64_t H, int64_t W,
                                         int64_t Cin, int64_t Cout, int64_t n_ctas, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(x && dy && partials, -1, "conv3x3_wgrad: null pointer argument");
    B200RL_REQUIRE(N >= 1 && H >= 1 && N * H < (1ll << 30), -2, "conv3x3_wgrad: N=%lld H=%lld", (long long)N, (long long)H);
    B200RL_REQUIRE(W == 8 || W == 16 || W == 32 || W == 64, -2, "conv3x3_wgrad: W=%lld (supported: 8, 16, 32, 64)", (long long)W);
    B200RL_REQUIRE((Cin == 3 || Cin == 16 || Cin == 32) && (Cout == 16 || Cout == 32), -2,
                   "conv3x3_wgrad: Cin=%lld Cout=%lld (supported: Cin 3/16/32, Cout 16/32)", (long long)Cin, (long long)Cout);
    B200RL_REQUIRE(n_ctas >= 1 && n_ctas <= 65535, -2, "conv3x3_wgrad: n_ctas=%lld", (long long)n_ctas);
    B200RL_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15u) == 0 && (reinterpret_cast<uintptr_t>(dy) & 15u) == 0, -3,
                   "conv3x3_wgrad: tensors must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    const int n = (int)N, h = (int)H, w = (int)W, ci = (int)Cin, g = (int)n_ctas;
    if (Cin == 3 && Cout == 16) return launch_wgrad_w<8, 16>(x, dy, partials, n, h, w, ci, g, s);
    if (Cin == 3 && Cout == 32) return launch_wgrad_w<8, 32>(x, dy, partials, n, h, w, ci, g, s);
    if (Cin == 16 && Cout == 16) return launch_wgrad_w<16, 16>(x, dy, partials, n, h, w, ci, g, s);
    if (Cin == 16 && Cout == 32) return launch_wgrad_w<16, 32>(x, dy, partials, n, h, w, ci, g, s);
    if (Cin == 32 && Cout == 16) return launch_wgrad_w<32, 16>(x, dy, partials, n, h, w, ci, g, s);
    return launch_wgrad_w<32, 32>(x, dy, partials, n, h, w, ci, g, s);
}

// =====================================================================================================================
// K11: the 3x3 / stride 1 / padding 1 convolution itself (forward, and -- with the weights read transposed and
// flipped -- the input gradient), bias-free, fp32 NHWC, TF32 tensor-core products with fp32 accumulation:
//
//   y[n,h,w,o] = sum over (kh, kw, i) of x[n, h+kh-1, w+kw-1, i] * w[o][kh][kw][i]                      (forward)
//   dx[n,h,w,i] = sum over (kh, kw, o) of dy[n, h-kh+1, w-kw+1, o] * w[o][kh][kw][i]                     (dgrad)
//     = the forward form with in/out channels swapped and weight w'[i][kh'][kw'][o] = w[o][2-kh'][2-kw'][i].
//
// Implicit GEMM with M = pixels, N = output channels, K = 9 x input channels.  A CTA walks tiles of 8 image rows
// (persistent; the x tile with its one-pixel halo arrives by cp.async, double-buffered); the whole weight tensor sits
// in shared memory for the CTA's lifetime; a warp owns W/16 (W = 8: one) m-tiles of 16 pixels and all output
// channels: per tap and 8 input channels it loads its A fragments (x, pixel stride = 4 mod 8 floats: conflict-free)
// and the B fragments (weights, row stride = 8 mod 32) and issues MT x Cout/8 MMAs.  Results go straight from the
// accumulator registers to global memory (8-byte stores, 32 contiguous bytes per pixel and n-tile).
// cuDNN picks sm80 "indexed_wo_smem" kernels for the 16-channel layers (1.06 ms per layer at 5,120 frames, 22 TFLOP/s).
namespace b200rl {

constexpr int kFwTH = 8;
// optional epilogue of the forward form: v = acc + bias[o] (+ res); y = v (when y != NULL); relu_out = max(v, 0)
struct FwEpilogue {
    const float *bias, *res;
    float *relu_out;
};
template <int W> struct FwWarps { static constexpr int value = W >= 16 ? 8 : 4; };

template <int CIN_P, int COUT, int W, bool DGRAD>
__global__ void __launch_bounds__(FwWarps<W>::value * 32)
conv3x3_fwd_kernel(const float *__restrict__ x, const float *__restrict__ wgt, float *__restrict__ y, FwEpilogue ep, int N,
                   int H, int Cin, int n_tiles) {
    constexpr int XS = CIN_P + 4;                          // pixel stride of the x tile: = 4 (mod 8)
    constexpr int WS = COUT + 8;                           // row stride of the weights [tap][cin][cout]: = 8 (mod 32)
    constexpr int WARPS = FwWarps<W>::value, THREADS = WARPS * 32;
    constexpr int XW = W + 2, x_floats = (kFwTH + 2) * XW * XS, w_floats = 9 * CIN_P * WS;
    constexpr int M_TILES = kFwTH * W / 16, MT = M_TILES / WARPS, NT = COUT / 8, KC = CIN_P / 8;
    constexpr int XBYTES = CIN_P == 8 ? 4 : 16;
    static_assert(M_TILES % WARPS == 0, "m-tiles must divide among the warps");
    extern __shared__ __align__(16) float smem[];
    float *ws = smem, *xbuf = smem + w_floats;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, gid = lane >> 2, tig = lane & 3;
    const int tiles_per_image = (H + kFwTH - 1) / kFwTH;
    const int x_per_row = W * Cin / (XBYTES / 4);
    const RowPlan xplan = make_plan<THREADS>(x_per_row, XBYTES / 4, Cin, XS, 1);

    // weights -> shared memory [tap][cin][cout] (zero rows for padded input channels); global layout [O][3][3][I]
    for (int i = tid; i < w_floats; i += THREADS) {
        const int o = i % WS, rest = i / WS, c = rest % CIN_P, tap = rest / CIN_P;
        float v = 0.f;
        if (o < COUT && c < Cin)
            v = DGRAD ? __ldg(wgt + ((long long)c * 9 + (8 - tap)) * COUT + o)       // w[o'=c][2-kh][2-kw][i'=o]
                      : __ldg(wgt + ((long long)o * 9 + tap) * Cin + c);
        ws[i] = tf32_rna(v);
    }
    for (int i = tid; i < 2 * x_floats; i += THREADS) xbuf[i] = 0.f;    // halo columns, padded channels: zero for good
    __syncthreads();

    auto stage = [&](int t, int b) {
        if (t < n_tiles) {
            const int n = t / tiles_per_image, h0 = (t - n * tiles_per_image) * kFwTH;
            stage_rows<XBYTES, THREADS>(xbuf + b * x_floats, XW * XS, x, (long long)W * Cin, (long long)n * H, h0 - 1, kFwTH + 2,
                                        H, xplan, x_per_row, Cin, XS, 1);
        }
        cp_async_commit();
    };

    // m-tile m of the tile: 16 consecutive pixels of a row (W >= 16) or 2 rows of 8 (W = 8); rows 0-7 of the fragment
    // are pixels +0..7, rows 8-15 are `up8` floats further
    constexpr int up8 = (W >= 16 ? 8 : XW) * XS;
    int tile = blockIdx.x, buf = 0;
    stage(tile, 0);
    for (; tile < n_tiles; tile += gridDim.x, buf ^= 1) {
        stage(tile + gridDim.x, buf ^ 1);
        cp_async_wait_group<1>();
        __syncthreads();
        const float *xs = xbuf + buf * x_floats;
        float acc[MT][NT][4];
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NT; ++b)
#pragma unroll
                for (int d = 0; d < 4; ++d) acc[a][b][d] = 0.f;
        int prow[MT], pcol[MT];
#pragma unroll
        for (int m = 0; m < MT; ++m) {
            const int mt = warp * MT + m;
            prow[m] = W >= 16 ? mt / (W / 16) : mt * 2;
            pcol[m] = W >= 16 ? (mt % (W / 16)) * 16 : 0;
        }
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const int kh = tap / 3, kw = tap % 3;
#pragma unroll
            for (int kc = 0; kc < KC; ++kc) {
                float bf[NT][2];
                const float *wp = ws + ((tap * CIN_P + kc * 8 + tig) * WS) + gid;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    bf[nt][0] = wp[nt * 8];
                    bf[nt][1] = wp[4 * WS + nt * 8];
                }
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                    const float *ap = xs + ((prow[m] + kh) * XW + pcol[m] + kw + gid) * XS + kc * 8 + tig;
                    float a[4];
                    a[0] = tf32_rna(ap[0]);
                    a[1] = tf32_rna(ap[up8]);
                    a[2] = tf32_rna(ap[4]);
                    a[3] = tf32_rna(ap[up8 + 4]);
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) mma_tf32_16x8x8(acc[m][nt], a, bf[nt][0], bf[nt][1]);
                }
            }
        }
        // ---- accumulators -> y (rows of the image that exist) ----
        const int n = tile / tiles_per_image, h0 = (tile - n * tiles_per_image) * kFwTH;
#pragma unroll
        for (int m = 0; m < MT; ++m) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int r = prow[m] + (W >= 16 ? 0 : half), c = pcol[m] + gid + (W >= 16 ? 8 * half : 0), h = h0 + r;
                if (h < H) {
                    const long long off = (((long long)n * H + h) * W + c) * COUT + 2 * tig;
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) {
                        float2 v = make_float2(acc[m][nt][2 * half], acc[m][nt][2 * half + 1]);
                        if (ep.bias) {                      // torch order: (conv + bias), then inputs + that
                            const float2 b = __ldg(reinterpret_cast<const float2 *>(ep.bias + nt * 8 + 2 * tig));
                            v.x += b.x; v.y += b.y;
                        }
                        if (ep.res) {
                            const float2 s2 = __ldcs(reinterpret_cast<const float2 *>(ep.res + off + nt * 8));
                            v.x = s2.x + v.x; v.y = s2.y + v.y;
                        }
                        if (y) __stcs(reinterpret_cast<float2 *>(y + off + nt * 8), v);
                        if (ep.relu_out)
                            __stcs(reinterpret_cast<float2 *>(ep.relu_out + off + nt * 8), make_float2(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f)));
                    }
                }
            }
        }
        __syncthreads();                                    // the buffer is refilled by the next round's copies
    }
    cp_async_wait_group<0>();
}

// ---- K11, row-streaming variant for the 16-output-channel layers at W >= 32 ------------------------------------------
// With 16 output channels an x fragment feeds only two MMAs above, and the weights are re-read from shared memory for
// every m-tile: 3 fragment words per MMA.  Here the whole weight tensor lives in REGISTERS (9 taps x Cin/8 x 2 n-tiles
// x 2 words = 72 at 16 -> 16 channels) and a warp walks down a strip of 16 columns: the x fragments of tile row j
// (column shift kw, 8 input channels) feed the output rows j, j-1, j-2 as kernel rows kh = 0, 1, 2 -- six MMAs per
// fragment, 0.67 words per MMA -- with three rolling accumulator rows; a row is stored as soon as its third x row is done.
template <int CIN_P, int COUT, int W>
struct Fw2 {
    static constexpr bool use = COUT == 16 && CIN_P <= 16 && W >= 32;
    static constexpr int RS = 4;                            // output rows per warp
    static constexpr int warps = (W / 16) * (kFwTH / RS);
};

template <int CIN_P, int COUT, int W, bool DGRAD>
__global__ void __launch_bounds__(Fw2<CIN_P, COUT, W>::warps * 32)
conv3x3_fwd_rows_kernel(const float *__restrict__ x, const float *__restrict__ wgt, float *__restrict__ y, FwEpilogue ep, int N,
                        int H, int Cin, int n_tiles) {
    constexpr int XS = CIN_P + 4;
    constexpr int WARPS = Fw2<CIN_P, COUT, W>::warps, THREADS = WARPS * 32, RS = Fw2<CIN_P, COUT, W>::RS;
    constexpr int XW = W + 2, x_floats = (kFwTH + 2) * XW * XS;
    constexpr int NT = COUT / 8, KC = CIN_P / 8;
    constexpr int XBYTES = CIN_P == 8 ? 4 : 16;
    extern __shared__ __align__(16) float smem[];
    float *xbuf = smem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, gid = lane >> 2, tig = lane & 3;
    const int strip = warp % (W / 16), seg = warp / (W / 16);
    const int tiles_per_image = (H + kFwTH - 1) / kFwTH;
    const int x_per_row = W * Cin / (XBYTES / 4);
    const RowPlan xplan = make_plan<THREADS>(x_per_row, XBYTES / 4, Cin, XS, 1);

    // this lane's B fragments of every tap: b0 = w'[tap][i = kc*8 + tig][o = nt*8 + gid], b1 = same at i + 4
    float bw[9][KC][NT][2];
#pragma unroll
    for (int tap = 0; tap < 9; ++tap)
#pragma unroll
        for (int kc = 0; kc < KC; ++kc)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    const int c = kc * 8 + tig + 4 * hf, o = nt * 8 + gid;
                    float v = 0.f;
                    if (c < Cin)
                        v = DGRAD ? __ldg(wgt + ((long long)c * 9 + (8 - tap)) * COUT + o)
                                  : __ldg(wgt + ((long long)o * 9 + tap) * Cin + c);
                    bw[tap][kc][nt][hf] = tf32_rna(v);
                }
    for (int i = tid; i < 2 * x_floats; i += THREADS) xbuf[i] = 0.f;
    __syncthreads();

    auto stage = [&](int t, int b) {
        if (t < n_tiles) {
            const int n = t / tiles_per_image, h0 = (t - n * tiles_per_image) * kFwTH;
            stage_rows<XBYTES, THREADS>(xbuf + b * x_floats, XW * XS, x, (long long)W * Cin, (long long)n * H, h0 - 1, kFwTH + 2,
                                        H, xplan, x_per_row, Cin, XS, 1);
        }
        cp_async_commit();
    };

    const int a_base = ((seg * RS) * XW + strip * 16 + gid) * XS + tig;   // tile row seg*RS, column strip*16 + gid (kw = 0)
    int tile = blockIdx.x, buf = 0;
    stage(tile, 0);
    for (; tile < n_tiles; tile += gridDim.x, buf ^= 1) {
        stage(tile + gridDim.x, buf ^ 1);
        cp_async_wait_group<1>();
        __syncthreads();
        const float *xs = xbuf + buf * x_floats + a_base;
        const int n = tile / tiles_per_image, h0 = (tile - n * tiles_per_image) * kFwTH;
        float acc[3][NT][4];
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int b = 0; b < NT; ++b)
#pragma unroll
                for (int d = 0; d < 4; ++d) acc[a][b][d] = 0.f;
#pragma unroll
        for (int j = 0; j < RS + 2; ++j) {                  // x tile row seg*RS + j
#pragma unroll
            for (int kw = 0; kw < 3; ++kw)
#pragma unroll
                for (int kc = 0; kc < KC; ++kc) {
                    const float *ap = xs + (j * XW + kw) * XS + kc * 8;
                    float a[4];
                    a[0] = tf32_rna(ap[0]);
                    a[1] = tf32_rna(ap[8 * XS]);
                    a[2] = tf32_rna(ap[4]);
                    a[3] = tf32_rna(ap[8 * XS + 4]);
#pragma unroll
                    for (int kh = 0; kh < 3; ++kh) {
                        const int i = j - kh;               // output row (local) that sees x row j as kernel row kh
                        if (i < 0 || i >= RS) continue;
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt)
                            mma_tf32_16x8x8(acc[i % 3][nt], a, bw[kh * 3 + kw][kc][nt][0], bw[kh * 3 + kw][kc][nt][1]);
                    }
                }
            if (j >= 2) {                                    // output row j - 2 has seen its three x rows: store, recycle
                const int i = j - 2, h = h0 + seg * RS + i;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    if (h < H) {
                        const long long off = (((long long)n * H + h) * W + strip * 16 + gid + 8 * half) * COUT + 2 * tig;
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt) {
                            float2 v = make_float2(acc[i % 3][nt][2 * half], acc[i % 3][nt][2 * half + 1]);
                            if (ep.bias) {
                                const float2 b = __ldg(reinterpret_cast<const float2 *>(ep.bias + nt * 8 + 2 * tig));
                                v.x += b.x; v.y += b.y;
                            }
                            if (ep.res) {
                                const float2 s2 = __ldcs(reinterpret_cast<const float2 *>(ep.res + off + nt * 8));
                                v.x = s2.x + v.x; v.y = s2.y + v.y;
                            }
                            if (y) __stcs(reinterpret_cast<float2 *>(y + off + nt * 8), v);
                            if (ep.relu_out)
                                __stcs(reinterpret_cast<float2 *>(ep.relu_out + off + nt * 8),
                                       make_float2(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f)));
                        }
                    }
                }
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                    for (int d = 0; d < 4; ++d) acc[i % 3][nt][d] = 0.f;
            }
        }
        __syncthreads();
    }
    cp_async_wait_group<0>();
}

template <typename Kern>
static int32_t launch_fwd_persistent(Kern kern, int threads, int smem, int *ctas_per_sm, const float *x, const float *wgt,
                                     float *y, FwEpilogue ep, int N, int H, int Cin, cudaStream_t stream) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 64) dev = 63;
    if (!ctas_per_sm[dev]) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
            return check_launch("conv3x3_fwd: shared-memory opt-in");
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, threads, smem) != cudaSuccess || nb < 1)
            return check_launch("conv3x3_fwd: occupancy query");
        ctas_per_sm[dev] = nb;
    }
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int n_tiles = N * ((H + kFwTH - 1) / kFwTH);
    int grid = sms * ctas_per_sm[dev];
    if (grid > n_tiles) grid = n_tiles;
    kern<<<grid, threads, smem, stream>>>(x, wgt, y, ep, N, H, Cin, n_tiles);
    return check_launch("conv3x3_fwd_kernel");
}

template <int CIN_P, int COUT, int W, bool DGRAD>
static int32_t launch_fwd(const float *x, const float *wgt, float *y, FwEpilogue ep, int N, int H, int Cin, cudaStream_t stream) {
    constexpr int XS = CIN_P + 4, WS = COUT + 8;
    static int ctas_per_sm[64] = {};
    if constexpr (Fw2<CIN_P, COUT, W>::use) {               // weights in registers: the tile buffers are all of it
        constexpr int smem = 4 * 2 * (kFwTH + 2) * (W + 2) * XS;
        return launch_fwd_persistent(conv3x3_fwd_rows_kernel<CIN_P, COUT, W, DGRAD>, Fw2<CIN_P, COUT, W>::warps * 32, smem,
                                     ctas_per_sm, x, wgt, y, ep, N, H, Cin, stream);
    } else {
        constexpr int smem = 4 * (9 * CIN_P * WS + 2 * (kFwTH + 2) * (W + 2) * XS);
        if (smem > 227 * 1024) {
            set_error("conv3x3_fwd: Cin=%d Cout=%d W=%d needs %d bytes of shared memory", Cin, COUT, W, smem);
            return -2;
        }
        return launch_fwd_persistent(conv3x3_fwd_kernel<CIN_P, COUT, W, DGRAD>, FwWarps<W>::value * 32, smem, ctas_per_sm, x,
                                     wgt, y, ep, N, H, Cin, stream);
    }
}

template <int CIN_P, int COUT, bool DGRAD>
static int32_t launch_fwd_w(const float *x, const float *wgt, float *y, FwEpilogue ep, int N, int H, int W, int Cin,
                            cudaStream_t stream) {
    switch (W) {
        case 8: return launch_fwd<CIN_P, COUT, 8, DGRAD>(x, wgt, y, ep, N, H, Cin, stream);
        case 16: return launch_fwd<CIN_P, COUT, 16, DGRAD>(x, wgt, y, ep, N, H, Cin, stream);
        case 32: return launch_fwd<CIN_P, COUT, 32, DGRAD>(x, wgt, y, ep, N, H, Cin, stream);
        case 64: return launch_fwd<CIN_P, COUT, 64, DGRAD>(x, wgt, y, ep, N, H, Cin, stream);
    }
    set_error("conv3x3_fwd: W=%d (supported: 8, 16, 32, 64)", W);
    return -2;
}

template <bool DGRAD>
static int32_t dispatch_fwd(const float *x, const float *wgt, float *y, FwEpilogue ep, int N, int H, int W, int Cin, int Cout,
                            cudaStream_t s) {
    if (Cin == 3 && Cout == 16) return launch_fwd_w<8, 16, DGRAD>(x, wgt, y, ep, N, H, W, Cin, s);
    if (Cin == 3 && Cout == 32) return launch_fwd_w<8, 32, DGRAD>(x, wgt, y, ep, N, H, W, Cin, s);
    if (Cin == 16 && Cout == 16) return launch_fwd_w<16, 16, DGRAD>(x, wgt, y, ep, N, H, W, Cin, s);
    if (Cin == 16 && Cout == 32) return launch_fwd_w<16, 32, DGRAD>(x, wgt, y, ep, N, H, W, Cin, s);
    if (Cin == 32 && Cout == 16) return launch_fwd_w<32, 16, DGRAD>(x, wgt, y, ep, N, H, W, Cin, s);
    return launch_fwd_w<32, 32, DGRAD>(x, wgt, y, ep, N, H, W, Cin, s);
}

}  // namespace b200rl

extern "C" int32_t b200rl_conv3x3_fwd(const float *x, const float *weight, float *y, int64_t N, int64_t H, int64_t W,
                                       int64_t Cin, int64_t Cout, int32_t dgrad, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(x && weight && y, -1, "conv3x3_fwd: null pointer argument");
    B200RL_REQUIRE(N >= 1 && H >= 1 && N * H < (1ll << 30), -2, "conv3x3_fwd: N=%lld H=%lld", (long long)N, (long long)H);
    B200RL_REQUIRE(W == 8 || W == 16 || W == 32 || W == 64, -2, "conv3x3_fwd: W=%lld (supported: 8, 16, 32, 64)", (long long)W);
    B200RL_REQUIRE((Cin == 3 || Cin == 16 || Cin == 32) && (Cout == 16 || Cout == 32), -2,
                   "conv3x3_fwd: Cin=%lld Cout=%lld (supported: in 3/16/32, out 16/32)", (long long)Cin, (long long)Cout);
    B200RL_REQUIRE(!(dgrad && Cin == 3), -2, "conv3x3_fwd: dgrad needs 16 or 32 input channels");
    B200RL_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15u) == 0 && (reinterpret_cast<uintptr_t>(y) & 15u) == 0, -3,
                   "conv3x3_fwd: tensors must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    const FwEpilogue none = {nullptr, nullptr, nullptr};
    return dgrad ? dispatch_fwd<true>(x, weight, y, none, (int)N, (int)H, (int)W, (int)Cin, (int)Cout, s)
                 : dispatch_fwd<false>(x, weight, y, none, (int)N, (int)H, (int)W, (int)Cin, (int)Cout, s);
}

extern "C" int32_t b200rl_conv3x3_fwd_fused(const float *x, const float *weight, const float *bias, const float *res, float *y,
                                             float *relu_out, int64_t N, int64_t H, int64_t W, int64_t Cin, int64_t Cout,
                                             void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(x && weight && bias && (y || relu_out), -1, "conv3x3_fwd_fused: null pointer argument");
    B200RL_REQUIRE(N >= 1 && H >= 1 && N * H < (1ll << 30), -2, "conv3x3_fwd_fused: N=%lld H=%lld", (long long)N, (long long)H);
    B200RL_REQUIRE(W == 8 || W == 16 || W == 32 || W == 64, -2, "conv3x3_fwd_fused: W=%lld (supported: 8, 16, 32, 64)", (long long)W);
    B200RL_REQUIRE((Cin == 3 || Cin == 16 || Cin == 32) && (Cout == 16 || Cout == 32), -2,
                   "conv3x3_fwd_fused: Cin=%lld Cout=%lld (supported: in 3/16/32, out 16/32)", (long long)Cin, (long long)Cout);
    const uintptr_t all = reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(res) |
                          reinterpret_cast<uintptr_t>(relu_out);
    B200RL_REQUIRE((all & 15u) == 0 && (reinterpret_cast<uintptr_t>(bias) & 7u) == 0, -3,
                   "conv3x3_fwd_fused: tensors must be 16-byte (bias: 8-byte) aligned");
    const FwEpilogue ep = {bias, res, relu_out};
    return dispatch_fwd<false>(x, weight, y, ep, (int)N, (int)H, (int)W, (int)Cin, (int)Cout, (cudaStream_t)stream);
}
