// Shared device/host helpers for the b200rl kernels (sm_100a only).
#pragma once
#include <cuda.h>  // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint, libcuda is not linked)
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/b200rl.h"

namespace b200rl {

// ---------------------------------------------------------------- host: errors
void set_error(const char *fmt, ...);
int32_t check_launch(const char *what);
bool encode_tmap_2d_f32(CUtensorMap *map, const void *base, uint64_t inner, uint64_t rows, uint64_t pitch_bytes,
                        uint32_t box_inner, uint32_t box_rows);  // api.cu

// Dynamic shared memory opt-in, cached per device (one process may drive several GPUs).
constexpr int kMaxDevices = 64;
template <typename F>
inline int32_t ensure_dynamic_smem(F func, size_t bytes, size_t (&cache)[kMaxDevices], const char *what) {
    if (bytes <= 48 * 1024) return 0;
    int dev = 0;
    cudaGetDevice(&dev);
    dev = (dev >= 0 && dev < kMaxDevices) ? dev : 0;
    if (bytes <= cache[dev]) return 0;
    cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) {
        set_error("%s: cudaFuncSetAttribute(%zu B shared): %s", what, bytes, cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    cache[dev] = bytes;
    return 0;
}

#define B200RL_REQUIRE(cond, code, ...)          \
    do {                                         \
        if (!(cond)) {                           \
            ::b200rl::set_error(__VA_ARGS__);    \
            return (code);                       \
        }                                        \
    } while (0)

constexpr int kNumScalars = 8;
// workspace layout: 256 B header (uint32 launch generation, uint32 poisoned; see grid_collect), then the
// published partial sums, uint64 words [8][kMaxPartialCtas] (launch tag << 32 | float bits), then the trace area.
constexpr int kMaxPartialCtas = 16384;
constexpr int64_t kTraceOffset = 256 + (int64_t)kMaxPartialCtas * kNumScalars * 8;
// B200RL_TRACE: [kMaxTraceCtas][16] uint64 per CTA: 8 %globaltimer stamps then 8 clock64 stamps (profiling aid)
constexpr int kMaxTraceCtas = 4096;
// K6 (clip + Adam): 64 B header (uint32 generation) + one published word per CTA
constexpr int64_t kOptOffset = kTraceOffset + (int64_t)kMaxTraceCtas * 16 * 8;
constexpr int kMaxOptCtas = 1024;
constexpr int64_t kWorkspaceBytes = kOptOffset + 64 + (int64_t)kMaxOptCtas * 8;

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
// L2 prefetch of a contiguous global range (a hint: no data reaches the SM).  16-byte aligned address and size.
__device__ __forceinline__ void bulk_prefetch_l2(const void *src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() {
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
// make generic-proxy smem writes visible to the async proxy (before bulk_s2g)
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// TMA 2-D tiled copies through a tensor map (SASS: UTMALDG / UTMASTG).  x = innermost coordinate (elements),
// y = row.  Shared-memory tiles are dense [box_rows][box_inner] and 128-byte aligned; elements of the box
// that fall outside the tensor are zero-filled on load and dropped on store.
__device__ __forceinline__ void tma_load_2d(void *dst_smem, const CUtensorMap *map, int x, int y, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::
            "r"(smem_u32(dst_smem)),
        "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap *map, const void *src_smem, int x, int y) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(map), "r"(x),
                 "r"(y), "r"(smem_u32(src_smem))
                 : "memory");
}

// Thread-block clusters: rank of this CTA, cluster-wide barrier, and a 16-byte read of another CTA's shared
// memory (distributed shared memory: mapa translates a local shared address to the peer's window).
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float4 ld_dsmem_f4(const void *local_smem, uint32_t rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_smem)), "r"(rank));
    float4 v;
    asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "r"(remote)
                 : "memory");
    return v;
}

// 16-byte cp.async (LDGSTS), L1-bypassing; src_bytes < 16 zero-fills the rest (0: nothing is read)
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem)), "l"(gmem), "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
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

// Programmatic dependent launch (PDL).  Both are no-ops for a launch without the attribute.
// pdl_launch_dependents: the next kernel in the stream may start being scheduled;
// pdl_wait: blocks until every kernel this launch depends on has completed and flushed.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Deterministic CTA-wide sum of NV per-thread values.  `red` is shared scratch of
// NV * blockDim.x floats.  Fixed summation order; totals are valid in warp 0 (all lanes).
// Costs NV stores per thread plus one warp's column sums -- far fewer issue slots than a
// shuffle tree in every warp.  All threads must call; ends with warp 0 only holding totals.
template <int NV>
__device__ __forceinline__ void cta_sum(const float (&vals)[NV], float *red, float (&tot)[NV]) {
    const int nthr = blockDim.x;
#pragma unroll
    for (int k = 0; k < NV; ++k) red[k * nthr + threadIdx.x] = vals[k];
    __syncthreads();
    if (threadIdx.x < 32) {
        // NV independent accumulators: all loads of a row of 32 are in flight together
#pragma unroll
        for (int k = 0; k < NV; ++k) tot[k] = 0.f;
#pragma unroll 2
        for (int i = threadIdx.x; i < nthr; i += 32) {
#pragma unroll
            for (int k = 0; k < NV; ++k) tot[k] += red[k * nthr + i];
        }
        // NV interleaved shuffle trees (step outermost: the NV chains hide each other's latency)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int k = 0; k < NV; ++k) tot[k] += __shfl_xor_sync(0xffffffffu, tot[k], o);
        }
    }
}

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ unsigned int ld_volatile_u32(const unsigned int *p) {
    unsigned int v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u32(unsigned int *p, unsigned int v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Deterministic grid-wide sum of NV per-CTA partial sums, without fences or atomics.
//
// Workspace header (the first 256 bytes): uint32 generation, uint32 poisoned.  Every CTA PUBLISHES each
// partial as one 64-bit word (launch tag << 32 | float bits) with a single store: the tag travels with
// the data, so no fence is needed and the CTA retires immediately.  The tag is the workspace's launch
// generation + 1 (never 0): words are never cleared, a word of an older launch can not be mistaken for
// this launch's (the 31-bit tag wraps after 2^31 launches on one workspace: a stale word could alias only if
// its CTA index had gone unused for exactly that many launches and the grid then grew back).  One collector CTA (the highest blockIdx: normally dispatched last, so it rarely waits)
// polls the words until every one carries this launch's tag, sums them in a fixed order --
// deterministic, bit-reproducible scalars -- hands the totals to `finalize(sums)` (one thread) and
// bumps the generation.
//
// Forward progress does NOT depend on the dispatch order: the polling CTA holds one CTA slot of one SM;
// every other CTA of the grid only needs some slot somewhere on the GPU and never waits for anything,
// so all of them publish eventually whatever order the hardware picks.  Only the polling time depends
// on the order.  A collector that still sees a missing word after kSpinLimitCycles (~1 s: the GPU is
// wedged by something else) reports NaN sums and POISONS the workspace: every later launch on it
// reports NaN too, until the owner zeroes the workspace again -- a stale word can therefore never be
// summed silently.
//
// The generation is read after `before_publish()` (the PDL wait): the previous launch on this
// workspace has completed by then, its collector included.
constexpr long long kSpinLimitCycles = 1ll << 31;

// tot[k]: this CTA's partial sums, valid in thread 0..NV-1 at least (cta_sum leaves them in all of warp 0).
// `red`: shared scratch of NV * blockDim.x floats.  All threads of the CTA must call.
__device__ __forceinline__ unsigned int grid_generation(const void *workspace) {
    return ld_volatile_u32(reinterpret_cast<const unsigned int *>(workspace));
}

template <int NV, typename Finalize>
__device__ __forceinline__ void grid_collect(const float (&tot)[NV], float *red, void *workspace,
                                             unsigned int gen, Finalize finalize) {
    static_assert(NV <= kNumScalars, "partials rows");
    unsigned int *hdr = reinterpret_cast<unsigned int *>(workspace);
    unsigned long long *words =
        reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(workspace) + 256);
    const unsigned int tag = (gen & 0x7fffffffu) + 1u;
    const unsigned int n_other = gridDim.x - 1;
    if (blockIdx.x != n_other) {
#pragma unroll
        for (int k = 0; k < NV; ++k)
            if (threadIdx.x == k)
                st_volatile_u64(&words[k * kMaxPartialCtas + blockIdx.x],
                                ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(tot[k]));
        return;
    }
    // collector
    constexpr int U = 4;
    float acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k] = 0.f;
    const long long t0 = clock64();
    bool timed_out = false;
    for (unsigned int c0 = threadIdx.x; c0 < n_other; c0 += U * blockDim.x) {
        unsigned long long w[U][NV];
        bool all;
        do {
            all = true;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const unsigned int c = c0 + u * blockDim.x;
#pragma unroll
                for (int k = 0; k < NV; ++k) {
                    w[u][k] = c < n_other ? ld_volatile_u64(&words[k * kMaxPartialCtas + c]) : ((unsigned long long)tag << 32);
                    all = all && (unsigned int)(w[u][k] >> 32) == tag;
                }
            }
        } while (!all && clock64() - t0 < kSpinLimitCycles);
        timed_out = timed_out || !all;
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int k = 0; k < NV; ++k) acc[k] += __uint_as_float((unsigned int)w[u][k]);
    }
    if (timed_out) st_volatile_u32(hdr + 1, 1u);
    float own[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) own[k] = tot[k];
    __syncthreads();  // `red` may still be read by warp 0 (the caller's cta_sum); the poison store is ordered too
    float all_tot[NV];
    cta_sum<NV>(acc, red, all_tot);
    if (threadIdx.x == 0) {
        const bool poisoned = ld_volatile_u32(hdr + 1) != 0u;
#pragma unroll
        for (int k = 0; k < NV; ++k) all_tot[k] = poisoned ? __int_as_float(0x7fc00000) : all_tot[k] + own[k];
        finalize(all_tot);
        st_volatile_u32(hdr, gen + 1u);
    }
}

// Grid-wide sum of NV per-thread values: CTA reduction (cta_sum) then grid_collect.
// `before_publish()` runs in every thread right before the workspace is touched (PDL wait).
template <int NV, typename Finalize, typename BeforePublish>
__device__ __forceinline__ void grid_reduce(float (&vals)[NV], float *red, void *workspace,
                                            Finalize finalize, BeforePublish before_publish) {
    float tot[NV];
    cta_sum<NV>(vals, red, tot);
    before_publish();
    grid_collect<NV>(tot, red, workspace, grid_generation(workspace), finalize);
}

// The same protocol for small persistent grids (gridDim.x <= a few hundred), run by ONE WARP of each CTA:
// no block-wide barrier anywhere.  tot[k] must be valid in every lane of the calling warp.  The collector
// warp's lanes take CTAs lane, lane + 32, ... and the partials are summed in a fixed order.
// warp_publish: a non-collector CTA's partials (any CTA but the last); warp_collect: the last CTA's duty.
template <int NV>
__device__ __forceinline__ void warp_publish(const float (&tot)[NV], void *workspace, unsigned int gen) {
    static_assert(NV <= kNumScalars, "partials rows");
    unsigned long long *words =
        reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(workspace) + 256);
    const unsigned int tag = (gen & 0x7fffffffu) + 1u, lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 0; k < NV; ++k)
        if (lane == k)
            st_volatile_u64(&words[k * kMaxPartialCtas + blockIdx.x],
                            ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(tot[k]));
}

template <int NV, typename Finalize>
__device__ __forceinline__ void warp_collect(const float (&tot)[NV], void *workspace, unsigned int gen,
                                             Finalize finalize) {
    static_assert(NV <= kNumScalars, "partials rows");
    unsigned int *hdr = reinterpret_cast<unsigned int *>(workspace);
    unsigned long long *words =
        reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(workspace) + 256);
    const unsigned int tag = (gen & 0x7fffffffu) + 1u;
    const unsigned int n_other = gridDim.x - 1, lane = threadIdx.x & 31;
    float acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k] = 0.f;
    const long long t0 = clock64();
    bool timed_out = false;
    constexpr int U = 5;  // CTAs per lane per batch: all U * NV loads are in flight together (160 CTAs = one batch)
    for (unsigned int c0 = lane; c0 < n_other; c0 += 32 * U) {
        unsigned long long w[U][NV];
        bool all;
        do {
            all = true;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const unsigned int c = c0 + 32 * u;
#pragma unroll
                for (int k = 0; k < NV; ++k) {
                    w[u][k] = c < n_other ? ld_volatile_u64(&words[k * kMaxPartialCtas + c]) : ((unsigned long long)tag << 32);
                    all = all && (unsigned int)(w[u][k] >> 32) == tag;
                }
            }
        } while (!all && clock64() - t0 < kSpinLimitCycles);
        timed_out = timed_out || !all;
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int k = 0; k < NV; ++k) acc[k] += __uint_as_float((unsigned int)w[u][k]);
    }
    timed_out = __any_sync(0xffffffffu, timed_out);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int k = 0; k < NV; ++k) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
    }
    if (lane == 0) {
        if (timed_out) st_volatile_u32(hdr + 1, 1u);
        const bool poisoned = timed_out || ld_volatile_u32(hdr + 1) != 0u;
#pragma unroll
        for (int k = 0; k < NV; ++k) acc[k] = poisoned ? __int_as_float(0x7fc00000) : acc[k] + tot[k];
        finalize(acc);
        st_volatile_u32(hdr, gen + 1u);
    }
}

}  // namespace b200rl
