// Micro-benchmark (round 2, K1 design): what does ONE small streaming launch cost on a B200?
// A kernel that only reads R bytes and writes Wb bytes (K1's traffic at T=20, B=4096, A=15: 11.47 MB in, 5.24 MB
// out), as a dependent chain of launches in a CUDA graph over rotating cold buffers (> L2), with and without PDL.
// This is the floor any single-launch K1 can reach: launch gap + first-byte latency + bytes / bandwidth.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/micro/copy_floor scripts/micro/copy_floor.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <algorithm>
#include <vector>

template <int UNROLL>
__global__ void __launch_bounds__(1024) stream_kernel(const float4 *__restrict__ in, float4 *__restrict__ out, int n_in, int n_out, int pdl) {
    if (pdl) {
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        asm volatile("griddepcontrol.wait;" ::: "memory");
    }
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = tid; k < n_in; k += nthr * UNROLL) {
        float4 v[UNROLL];
#pragma unroll
        for (int j = 0; j < UNROLL; ++j)
            if (k + j * nthr < n_in) v[j] = __ldcs(in + k + j * nthr);
#pragma unroll
        for (int j = 0; j < UNROLL; ++j)
            if (k + j * nthr < n_in) {
                acc.x += v[j].x; acc.y += v[j].y; acc.z += v[j].z; acc.w += v[j].w;
                const int o = k + j * nthr;
                if (o < n_out) __stcs(out + o, v[j]);
            }
    }
    if (acc.x == 1.2345f) out[0] = acc;  // keep the loads
}

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

// The same traffic through the TMA unit: every CTA bulk-loads its share in `nch` chunks (one mbarrier each, all issued
// at once by `nissue` different warps) and bulk-stores the leading Wb/R part of every chunk as soon as it has landed.
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

__global__ void __launch_bounds__(640) tma_stream_kernel(const char *__restrict__ in, char *__restrict__ out, int r_cta, int w_cta, int nch, int pdl,
                                                          unsigned long long *stamps) {
    extern __shared__ __align__(128) unsigned char sm[];
    __shared__ unsigned long long bars[64];
    if (threadIdx.x == 0) {
        for (int k = 0; k < nch; ++k) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[k])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (pdl) {
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        asm volatile("griddepcontrol.wait;" ::: "memory");
    }
    __syncthreads();
    if (stamps && threadIdx.x == 0) stamps[blockIdx.x * 4 + 0] = gtime();
    const int rc = r_cta / nch, wc = w_cta / nch;   // multiples of 16 by construction
    const char *src = in + (size_t)blockIdx.x * r_cta;
    char *dst = out + (size_t)blockIdx.x * w_cta;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    if (lane == 0)
        for (int k = warp; k < nch; k += nw) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[k])), "r"(rc) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                             "r"(smem_u32(sm + (size_t)k * rc)), "l"(src + (size_t)k * rc), "r"(rc), "r"(smem_u32(&bars[k])) : "memory");
        }
    if (lane == 0)
        for (int k = warp; k < nch; k += nw) {
            unsigned done = 0;
            while (!done)
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(done) : "r"(smem_u32(&bars[k])) : "memory");
            if (stamps && k == 0) stamps[blockIdx.x * 4 + 1] = gtime();
            if (stamps && k == nch - 1) stamps[blockIdx.x * 4 + 2] = gtime();
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + (size_t)k * wc), "r"(smem_u32(sm + (size_t)k * rc)), "r"(wc) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    __syncthreads();
    if (stamps && threadIdx.x == 0) stamps[blockIdx.x * 4 + 3] = gtime();
}

int main() {
    const size_t R = 11468800, Wb = 5242880;
    const int nsets = 24, chain = 100, reps = 20;
    char *in, *out;
    cudaMalloc(&in, R * nsets);
    cudaMalloc(&out, Wb * nsets);
    cudaMemset(in, 0, R * nsets);
    cudaStream_t s;
    cudaStreamCreate(&s);
    printf("%6s %6s %4s %4s | %9s %9s\n", "ctas", "thr", "unr", "pdl", "us/launch", "GB/s");
    for (int pdl = 0; pdl < 2; ++pdl)
        for (int unr : {4, 8})
            for (int thr : {256, 512, 1024})
                for (int cps : {1, 2, 4}) {
                    if (thr * cps > 2048) continue;
                    const int grid = 148 * cps;
                    cudaGraph_t g;
                    cudaGraphExec_t ge;
                    cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
                    for (int i = 0; i < chain; ++i) {
                        cudaLaunchConfig_t cfg = {};
                        cfg.gridDim = dim3(grid);
                        cfg.blockDim = dim3(thr);
                        cfg.stream = s;
                        cudaLaunchAttribute at[1];
                        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
                        at[0].val.programmaticStreamSerializationAllowed = 1;
                        cfg.attrs = at;
                        cfg.numAttrs = pdl;
                        const float4 *ip = (const float4 *)(in + (size_t)(i % nsets) * R);
                        float4 *op = (float4 *)(out + (size_t)(i % nsets) * Wb);
                        if (unr == 4) cudaLaunchKernelEx(&cfg, stream_kernel<4>, ip, op, (int)(R / 16), (int)(Wb / 16), pdl);
                        else cudaLaunchKernelEx(&cfg, stream_kernel<8>, ip, op, (int)(R / 16), (int)(Wb / 16), pdl);
                    }
                    cudaStreamEndCapture(s, &g);
                    cudaGraphInstantiate(&ge, g, 0);
                    cudaGraphLaunch(ge, s);
                    cudaStreamSynchronize(s);
                    std::vector<float> ts;
                    for (int r = 0; r < reps; ++r) {
                        cudaEvent_t e0, e1;
                        cudaEventCreate(&e0); cudaEventCreate(&e1);
                        cudaEventRecord(e0, s);
                        cudaGraphLaunch(ge, s);
                        cudaEventRecord(e1, s);
                        cudaEventSynchronize(e1);
                        float ms; cudaEventElapsedTime(&ms, e0, e1);
                        ts.push_back(ms * 1e3f / chain);
                        cudaEventDestroy(e0); cudaEventDestroy(e1);
                    }
                    std::sort(ts.begin(), ts.end());
                    const float us = ts[ts.size() / 2];
                    printf("%6d %6d %4d %4d | %9.3f %9.0f\n", grid, thr, unr, pdl, us, (R + Wb) / us / 1e3);
                    cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
                }
    unsigned long long *stamps;
    cudaMalloc(&stamps, 8 * 4 * 148 * 4);
    cudaFuncSetAttribute(tma_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    printf("TMA variant: %6s %4s %4s | %9s %9s\n", "ctas", "nch", "pdl", "us/launch", "GB/s");
    for (int pdl = 0; pdl < 2; ++pdl)
        for (int cps : {1, 2})
            for (int nch : {1, 2, 4, 8, 16}) {
                const int grid = 148 * cps;
                // per-CTA shares, rounded down to multiples of 16 * nch (the few bytes lost do not matter here)
                const int r_cta = (int)(R / grid) / (16 * nch) * (16 * nch), w_cta = (int)(Wb / grid) / (16 * nch) * (16 * nch);
                cudaGraph_t g;
                cudaGraphExec_t ge;
                cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
                for (int i = 0; i < chain; ++i) {
                    cudaLaunchConfig_t cfg = {};
                    cfg.gridDim = dim3(grid);
                    cfg.blockDim = dim3(256);
                    cfg.dynamicSmemBytes = r_cta + 128;
                    cfg.stream = s;
                    cudaLaunchAttribute at[1];
                    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
                    at[0].val.programmaticStreamSerializationAllowed = 1;
                    cfg.attrs = at;
                    cfg.numAttrs = pdl;
                    cudaLaunchKernelEx(&cfg, tma_stream_kernel, (const char *)(in + (size_t)(i % nsets) * R), out + (size_t)(i % nsets) * Wb, r_cta, w_cta, nch, pdl,
                                       (unsigned long long *)(i == chain - 1 ? stamps : nullptr));
                }
                cudaStreamEndCapture(s, &g);
                cudaGraphInstantiate(&ge, g, 0);
                cudaGraphLaunch(ge, s);
                cudaStreamSynchronize(s);
                std::vector<float> ts;
                for (int r = 0; r < reps; ++r) {
                    cudaEvent_t e0, e1;
                    cudaEventCreate(&e0); cudaEventCreate(&e1);
                    cudaEventRecord(e0, s);
                    cudaGraphLaunch(ge, s);
                    cudaEventRecord(e1, s);
                    cudaEventSynchronize(e1);
                    float ms; cudaEventElapsedTime(&ms, e0, e1);
                    ts.push_back(ms * 1e3f / chain);
                    cudaEventDestroy(e0); cudaEventDestroy(e1);
                }
                std::sort(ts.begin(), ts.end());
                const float us = ts[ts.size() / 2];
                {
                    std::vector<unsigned long long> h(grid * 4);
                    cudaMemcpy(h.data(), stamps, h.size() * 8, cudaMemcpyDeviceToHost);
                    unsigned long long t0 = ~0ull;
                    for (int c = 0; c < grid; ++c) t0 = std::min(t0, h[c * 4]);
                    std::vector<double> a, b, d;
                    for (int c = 0; c < grid; ++c) { a.push_back(h[c * 4 + 1] - t0); b.push_back(h[c * 4 + 2] - t0); d.push_back(h[c * 4 + 3] - t0); }
                    std::sort(a.begin(), a.end()); std::sort(b.begin(), b.end()); std::sort(d.begin(), d.end());
                    printf("TMA variant: %6d %4d %4d | %9.3f %9.0f | ns since first CTA start: first chunk landed med %.0f max %.0f, last chunk med %.0f max %.0f, end max %.0f\n",
                           grid, nch, pdl, us, (R + Wb) / us / 1e3, a[grid / 2], a.back(), b[grid / 2], b.back(), d.back());
                }
                cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
            }
    printf("CTA footprint (148 CTAs, nch 8, pdl 1): %6s %8s | %9s\n", "thr", "smemKB", "us/launch");
    for (int thr : {256, 640})
        for (int extra : {0, 100 * 1024}) {
            const int grid = 148, nch = 8, pdl = 1;
            const int r_cta = (int)(R / grid) / (16 * nch) * (16 * nch), w_cta = (int)(Wb / grid) / (16 * nch) * (16 * nch);
            cudaFuncSetAttribute(tma_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            cudaGraph_t g;
            cudaGraphExec_t ge;
            cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
            for (int i = 0; i < chain; ++i) {
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3(grid);
                cfg.blockDim = dim3(thr);
                cfg.dynamicSmemBytes = r_cta + 128 + extra;
                cfg.stream = s;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
                at[0].val.programmaticStreamSerializationAllowed = 1;
                cfg.attrs = at;
                cfg.numAttrs = pdl;
                cudaLaunchKernelEx(&cfg, tma_stream_kernel, (const char *)(in + (size_t)(i % nsets) * R), out + (size_t)(i % nsets) * Wb, r_cta, w_cta, nch, pdl,
                                   (unsigned long long *)nullptr);
            }
            cudaStreamEndCapture(s, &g);
            cudaGraphInstantiate(&ge, g, 0);
            cudaGraphLaunch(ge, s);
            cudaStreamSynchronize(s);
            std::vector<float> ts;
            for (int r = 0; r < reps; ++r) {
                cudaEvent_t e0, e1;
                cudaEventCreate(&e0); cudaEventCreate(&e1);
                cudaEventRecord(e0, s);
                cudaGraphLaunch(ge, s);
                cudaEventRecord(e1, s);
                cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                ts.push_back(ms * 1e3f / chain);
                cudaEventDestroy(e0); cudaEventDestroy(e1);
            }
            std::sort(ts.begin(), ts.end());
            printf("CTA footprint: %6d %8d | %9.3f\n", thr, (r_cta + extra) / 1024, ts[ts.size() / 2]);
            cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
        }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
