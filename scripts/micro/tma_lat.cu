// Micro-benchmark (round 2, K1 design): latency and completion ORDER of TMA 1-D bulk loads (cp.async.bulk, UBLKCP).
// One CTA per SM; thread 0 issues N copies of SZ bytes each, every copy on its own mbarrier, then polls the
// barriers in issue order and stamps clock64 when each one is seen complete.  Cold data (a fresh region per trial).
// Prints, per (N, SZ): issue time of all N copies, completion time of the first / median / last copy (cycles,
// median over CTAs) and the achieved chip bandwidth.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/micro/tma_lat scripts/micro/tma_lat.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <algorithm>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int MAXN = 64;

__global__ void __launch_bounds__(128, 1) tma_kernel(const char *src, size_t cta_stride, int n, int sz, int use_ldg,
                                                      long long *out) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bars[MAXN];
    __shared__ long long stamps[MAXN + 2];
    const char *base = src + (size_t)blockIdx.x * cta_stride;
    if (threadIdx.x == 0) {
        for (int k = 0; k < n; ++k)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[k])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (use_ldg) {
        // reference: the same bytes by 128 threads of 16-byte loads, all issued before the first use
        const long long t0 = clock64();
        float4 acc = make_float4(0, 0, 0, 0);
        const int total = n * sz / 16;
        for (int k = threadIdx.x; k < total; k += 128 * 8) {
            float4 v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (k + j * 128 < total) v[j] = __ldcs(reinterpret_cast<const float4 *>(base) + k + j * 128);
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (k + j * 128 < total) { acc.x += v[j].x; acc.y += v[j].y; acc.z += v[j].z; acc.w += v[j].w; }
        }
        reinterpret_cast<float4 *>(smem)[threadIdx.x] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            const long long t1 = clock64();
            long long *o = out + (size_t)blockIdx.x * (MAXN + 2);
            o[0] = 0;
            for (int k = 0; k < n; ++k) o[1 + k] = t1 - t0;
        }
        return;
    }
    if (threadIdx.x == 0) {
        const long long t0 = clock64();
        for (int k = 0; k < n; ++k) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[k])), "r"(sz) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                             "r"(smem_u32(smem + (size_t)k * sz)), "l"(base + (size_t)k * sz), "r"(sz), "r"(smem_u32(&bars[k])) : "memory");
        }
        stamps[0] = clock64() - t0;
        for (int k = 0; k < n; ++k) {
            uint32_t done = 0;
            while (!done)
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(done) : "r"(smem_u32(&bars[k])) : "memory");
            stamps[1 + k] = clock64() - t0;
        }
        long long *o = out + (size_t)blockIdx.x * (MAXN + 2);
        for (int k = 0; k <= n; ++k) o[k] = stamps[k];
    }
}

int main() {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    const size_t region = (size_t)sms * 160 * 1024;           // one trial's footprint (<= 160 KB per CTA)
    const int trials = 12;
    char *buf;
    cudaMalloc(&buf, region * trials * 2);                     // > L2
    cudaMemset(buf, 1, region * trials * 2);
    long long *out;
    cudaMalloc(&out, sizeof(long long) * sms * (MAXN + 2));
    cudaFuncSetAttribute(tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    printf("SMs %d, clock %.0f MHz (nominal)\n", sms, khz / 1e3);
    printf("%5s %7s %4s | %8s | %8s %8s %8s | %9s\n", "N", "SZ", "ldg", "issue", "first", "median", "last", "GB/s@last");
    const int szs[] = {256, 1024, 4096, 16384};
    for (int use_ldg = 0; use_ldg < 2; ++use_ldg)
        for (int sz : szs)
            for (int n : {1, 2, 4, 8, 16, 32, 64}) {
                if ((size_t)n * sz > 160 * 1024) continue;
                std::vector<double> issue, first, med, last;
                for (int t = 0; t < trials; ++t) {
                    tma_kernel<<<sms, 128, 160 * 1024 + 256>>>(buf + (size_t)(t + (use_ldg ? trials : 0)) * region, 160 * 1024, n, sz, use_ldg, out);
                    std::vector<long long> h((size_t)sms * (MAXN + 2));
                    cudaMemcpy(h.data(), out, h.size() * 8, cudaMemcpyDeviceToHost);
                    if (t < 2) continue;
                    std::vector<double> a, b, c, d;
                    for (int s = 0; s < sms; ++s) {
                        long long *o = &h[(size_t)s * (MAXN + 2)];
                        a.push_back(o[0]); b.push_back(o[1]); c.push_back(o[1 + n / 2]); d.push_back(o[n]);
                    }
                    auto m = [](std::vector<double> &v) { std::sort(v.begin(), v.end()); return v[v.size() / 2]; };
                    issue.push_back(m(a)); first.push_back(m(b)); med.push_back(m(c)); last.push_back(m(d));
                }
                auto m = [](std::vector<double> &v) { std::sort(v.begin(), v.end()); return v[v.size() / 2]; };
                const double tl = m(last);
                printf("%5d %7d %4d | %8.0f | %8.0f %8.0f %8.0f | %9.0f\n", n, sz, use_ldg, m(issue), m(first), m(med), tl,
                       (double)n * sz * sms / (tl / (1.965e9)) / 1e9);
            }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
