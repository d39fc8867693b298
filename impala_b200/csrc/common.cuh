// Shared device/host helpers for the b200rl kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/b200rl.h"

namespace b200rl {

// ---------------------------------------------------------------- host: errors
void set_error(const char *fmt, ...);
int32_t check_launch(const char *what);

#define B200RL_REQUIRE(cond, code, ...)          \
    do {                                         \
        if (!(cond)) {                           \
            ::b200rl::set_error(__VA_ARGS__);    \
            return (code);                       \
        }                                        \
    } while (0)

constexpr int kNumScalars = 8;
// workspace layout: [0] uint32 ticket counter (self-resetting), pad to 256 B, then
// per-CTA partial sums [kMaxPartialCtas][8] floats.
constexpr int kMaxPartialCtas = 16384;
constexpr int64_t kWorkspaceBytes = 256 + (int64_t)kMaxPartialCtas * kNumScalars * 4;

// ---------------------------------------------------------------- device: PTX
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    }
}
// TMA 1-D bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP).
// src, dst 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src, uint32_t bytes,
                                         uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(dst_smem)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
// TMA 1-D bulk copy shared -> global (bulk async-group completion).
__device__ __forceinline__ void bulk_s2g(void *dst, const void *src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst),
                 "r"(smem_u32(src_smem)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() {
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
// make generic-proxy smem writes visible to the async proxy (before bulk_s2g)
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ float ld_stream_f32(const float *p) {
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}

// ---------------------------------------------------------------- device: math
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_log2(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Deterministic grid-wide sum of NV per-thread values.  Every CTA writes its
// partial to workspace, takes a ticket; the last CTA sums all partials in a fixed
// order, hands them to `finalize(sums)` (thread 0) and resets the ticket.
// `red` is shared scratch of at least NV * 32 floats.  All threads must call.
template <int NV, typename Finalize>
__device__ __forceinline__ void grid_reduce(float (&vals)[NV], float *red, void *workspace,
                                            Finalize finalize) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    unsigned int *ticket = reinterpret_cast<unsigned int *>(workspace);
    float *partials = reinterpret_cast<float *>(reinterpret_cast<char *>(workspace) + 256);
    __shared__ bool is_last;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        float s = warp_sum(vals[k]);
        if (lane == 0) red[k * 32 + warp] = s;
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            float s = lane < nwarps ? red[k * 32 + lane] : 0.f;
            s = warp_sum(s);
            if (lane == 0) partials[(size_t)blockIdx.x * kNumScalars + k] = s;
        }
        if (lane == 0) {
            __threadfence();
            unsigned int t = atomicAdd(ticket, 1u);
            is_last = (t == gridDim.x - 1);
        }
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // fixed-order sum: thread i takes CTAs i, i+blockDim, ... then a block tree
    float acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k] = 0.f;
    for (unsigned int c = threadIdx.x; c < gridDim.x; c += blockDim.x) {
#pragma unroll
        for (int k = 0; k < NV; ++k) acc[k] += __ldcg(&partials[(size_t)c * kNumScalars + k]);
    }
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        float s = warp_sum(acc[k]);
        if (lane == 0) red[k * 32 + warp] = s;
    }
    __syncthreads();
    if (warp == 0) {
        float tot[NV];
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            float s = lane < nwarps ? red[k * 32 + lane] : 0.f;
            tot[k] = warp_sum(s);
        }
        if (lane == 0) {
            finalize(tot);
            *ticket = 0u;  // leave the workspace ready for the next launch
        }
    }
}

}  // namespace b200rl
