// N5: fused actor epilogue -- softmax -> categorical sample -> argmax select, one thread per row.
// Replaces reference AtariPPOModel.act (models/distributed_models.py:26-31):
//   action = logits.softmax(-1).multinomial(1);  action = where(deterministic, logits.argmax(-1), action)
// torch's multinomial stream cannot be reproduced, so the draw is specified like the replay sampler:
// row g of the call uses Philox4x32-10 counter (lo32(g'), hi32(g'), stream_id, 1), g' = draw_offset + row,
// u = (out0 >> 8) * 2^-24 in [0,1);  a = first j with cumsum_j(e) > u * sum(e),  e_j = exp(x_j - max x).
#include "common.cuh"

namespace b200rl {

__device__ __forceinline__ void philox4x32_10_act(uint32_t &c0, uint32_t &c1, uint32_t &c2, uint32_t &c3,
                                                  uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}

__global__ void __launch_bounds__(128) act_epilogue_kernel(const float *__restrict__ logits, long long n, int A,
                                                           const unsigned char *__restrict__ deterministic,
                                                           long long det_stride, uint64_t seed, uint32_t stream_id,
                                                           uint64_t draw_offset, long long *__restrict__ action) {
    const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    const float *x = logits + row * A;
    float m = -INFINITY;
    int arg = 0;
    for (int j = 0; j < A; ++j) {
        const float v = __ldg(x + j);
        if (v > m) { m = v; arg = j; }  // first maximum, as torch.argmax
    }
    int a = arg;
    if (!deterministic[row * det_stride]) {
        float total = 0.f;
        for (int j = 0; j < A; ++j) total += __expf(__ldg(x + j) - m);
        const uint64_t g = draw_offset + (uint64_t)row;
        uint32_t c0 = (uint32_t)g, c1 = (uint32_t)(g >> 32), c2 = stream_id, c3 = 1u;
        philox4x32_10_act(c0, c1, c2, c3, (uint32_t)seed, (uint32_t)(seed >> 32));
        const float target = (float)(c0 >> 8) * 0x1.0p-24f * total;
        float cum = 0.f;
        a = A - 1;
        for (int j = 0; j < A; ++j) {
            cum += __expf(__ldg(x + j) - m);
            if (cum > target) { a = j; break; }
        }
    }
    action[row] = a;
}

}  // namespace b200rl

extern "C" int32_t b200rl_act_epilogue(const float *logits, int64_t n, int64_t A, const uint8_t *deterministic,
                                       int64_t deterministic_stride, uint64_t seed, uint32_t stream_id,
                                       uint64_t draw_offset, int64_t *action, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(n >= 0, -2, "act_epilogue: n < 0");
    if (n == 0) return 0;
    B200RL_REQUIRE(logits && deterministic && action, -1, "act_epilogue: null pointer argument");
    B200RL_REQUIRE(A >= 1 && A < (1 << 20), -2, "act_epilogue: A=%lld", (long long)A);
    B200RL_REQUIRE(deterministic_stride == 0 || deterministic_stride == 1, -4, "act_epilogue: deterministic stride must be 0 (broadcast) or 1");
    act_epilogue_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        logits, n, (int)A, deterministic, deterministic_stride, seed, stream_id, draw_offset,
        reinterpret_cast<long long *>(action));
    return check_launch("act_epilogue_kernel");
}
