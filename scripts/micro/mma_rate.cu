// Micro-benchmark: issue rate of the legacy warp-level tensor path (mma.sync.m16n8k8 TF32) on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu && ./mma_rate
#include <cstdio>
#include <cuda_runtime.h>
template <int ACC>
__global__ void __launch_bounds__(768, 1) k(float *out, int iters) {
    float acc[ACC][4];
    for (int i = 0; i < ACC; ++i) for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    unsigned a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ACC; ++i)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(acc[i][0]), "+f"(acc[i][1]), "+f"(acc[i][2]), "+f"(acc[i][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    float s = 0.f;
    for (int i = 0; i < ACC; ++i) for (int j = 0; j < 4; ++j) s += acc[i][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ACC>
void run(int threads) {
    float *out; cudaMalloc(&out, 148 * 1024 * 4);
    const int iters = 20000;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<ACC><<<148, threads>>>(out, 100);
    cudaEventRecord(e0);
    k<ACC><<<148, threads>>>(out, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double mmas = 148.0 * (threads / 32) * iters * ACC;
    printf("warps/SM %2d, %2d independent accumulators: %.1f TFLOP/s TF32 (%.2f cycles/MMA/SM at 1.965 GHz)\n", threads / 32, ACC,
           mmas * 2048 / ms / 1e9, ms * 1e-3 * 1.965e9 / (mmas / 148));
    cudaFree(out);
}
int main() { run<1>(128); run<4>(128); run<12>(128); run<12>(256); run<12>(768); run<4>(768); run<1>(768); return 0; }
