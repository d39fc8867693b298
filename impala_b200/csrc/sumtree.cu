// N3: prioritized sampling on the device -- sum tree update + inverse-CDF sampling.
// Replaces rlmeta PrioritizedSampler as built at reference agents/dqn/builder.py:83-90 and used by
// ReplayBuffer.extend(items, priorities) / sample / update (agents/dqn/learning.py:54,122,184).
// rlmeta's C++ RNG stream is unobtainable (PARITY UNPINNED), so the sampler is specified in
// oracle/replay_ref.py and reproduced here bit for bit:
//   tree: complete binary tree of fp64 sums over P = 2^L leaves, node i = node 2i + node 2i+1
//         (one IEEE add per node: CPU and GPU build identical trees from identical leaves);
//   draw g of stream (seed, stream_id): x64 from Philox4x32-10 exactly as the uniform sampler;
//         mass = (x64 >> 11) * 2^-53 * total;  descend: left = tree[2i]; mass < left ? go left :
//         (mass -= left, go right);  index = min(leaf, size - 1);  probability = leaf value / total.
#include "common.cuh"

namespace b200rl {

__device__ __forceinline__ void philox4x32_10_st(uint32_t &c0, uint32_t &c1, uint32_t &c2, uint32_t &c3,
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

// Update = set n leaves, refresh their ancestors.  A learner step updates a few thousand leaves of a tree
// with ~10^6; one SM cannot issue the scattered loads fast enough (measured: 71 us for 4096 of 1M from one
// CTA), and a level-by-level walk is a chain of dependent L2 round trips.  So:
//   * the tree is cut into G subtrees (roots = nodes [G, 2G)); CTA c scans all n indices, keeps those that
//     fall under its root (n/G on average) and refreshes only its own subtree -- no cross-CTA ordering;
//   * inside a subtree the levels wider than kSub are walked level by level (children fetched with one
//     16-byte load); the kSub-wide level is then read ONCE into shared memory and everything above it is
//     reduced there (recomputing an untouched node reproduces its value bit for bit, since
//     node = left + right always holds);
//   * a second one-CTA launch reduces the top log2(G) levels the same way from one read of level G.
// Duplicate indices: one of their values wins (in the learner's use -- keys sampled with replacement --
// duplicates carry the same priority, so the choice is immaterial).
constexpr int kSub = 512;           // widest level of a subtree held in shared memory
constexpr int kUpdThreads = 256;
constexpr int kUpdMax = 8192;       // updates per launch of this path
constexpr int kMaxSubtrees = 128;   // <= kSub so that the top launch fits the same shared array

// Reduce the subtree rooted at `root` above its `sb`-wide level (sb <= kSub, a power of two; that level is
// final in global memory) in shared memory and write every node above that level back.
__device__ __forceinline__ void sumtree_reduce_top(double *tree, long long root, int sb, double *s_top) {
    const int tid = threadIdx.x;
    const long long first = root * sb;  // the sb-wide level of this subtree: nodes [root*sb, (root+1)*sb)
    for (int i = tid; i < sb; i += blockDim.x) s_top[sb + i] = __ldcg(tree + first + i);
    __syncthreads();
    for (int w = sb >> 1; w >= 1; w >>= 1) {
        for (int i = tid; i < w; i += blockDim.x) s_top[w + i] = s_top[2 * (w + i)] + s_top[2 * (w + i) + 1];
        __syncthreads();
    }
    for (int q = 1 + tid; q < sb; q += blockDim.x) {
        const int l = 31 - __clz(q);  // local level of local node q in [2^l, 2^(l+1))
        tree[(root << l) + (q - (1 << l))] = s_top[q];
    }
}

__global__ void __launch_bounds__(kUpdThreads) sumtree_update_kernel(double *tree, long long P, int G, const long long *idx,
                                                                     const double *value, int n) {
    __shared__ int s_leaf[kUpdMax];
    __shared__ double s_top[2 * kSub];
    __shared__ int s_count;
    const int tid = threadIdx.x;
    const long long S = P / G;              // leaves per subtree
    const long long lo = (long long)blockIdx.x * S;
    if (tid == 0) s_count = 0;
    __syncthreads();
    for (int i = tid; i < n; i += kUpdThreads) {
        const long long j = __ldg(idx + i);
        if (j >= lo && j < lo + S && j >= 0 && j < P) {
            s_leaf[atomicAdd(&s_count, 1)] = (int)j;
            __stcg(tree + P + j, value[i]);
        }
    }
    __syncthreads();
    const int m = s_count;
    const int sb = S < kSub ? (int)S : kSub;
    if (m == 0) return;                     // untouched subtree: every node already equals the sum of its children
    int shift = 1;
    for (long long width = S >> 1; width >= sb; width >>= 1, ++shift) {
        for (int i = tid; i < m; i += kUpdThreads) {
            const long long node = (P + s_leaf[i]) >> shift;
            const double2 ch = __ldcg(reinterpret_cast<const double2 *>(tree + 2 * node));
            __stcg(tree + node, ch.x + ch.y);  // duplicates recompute the same value
        }
        __syncthreads();
    }
    sumtree_reduce_top(tree, (long long)G + blockIdx.x, sb, s_top);
}

// Top of the tree: nodes [G, 2G) are final; reduce the log2(G) levels above them.
__global__ void __launch_bounds__(kUpdThreads) sumtree_top_kernel(double *tree, int G) {
    __shared__ double s_top[2 * kSub];
    sumtree_reduce_top(tree, 1, G, s_top);
}

// Full rebuild of one level from the level below (used for bulk loads): node in [width, 2*width).
__global__ void __launch_bounds__(256) sumtree_level_kernel(double *tree, long long width) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < width) tree[width + i] = tree[2 * (width + i)] + tree[2 * (width + i) + 1];
}

__global__ void __launch_bounds__(256) sumtree_set_leaves_kernel(double *tree, long long P, const long long *idx,
                                                                 const double *value, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long j = idx[i];
    if (j >= 0 && j < P) tree[P + j] = value[i];
}

__global__ void __launch_bounds__(256) sumtree_sample_kernel(const double *__restrict__ tree, long long P,
                                                             long long size, uint64_t seed, uint32_t stream_id,
                                                             uint64_t draw_offset, long long n,
                                                             long long *__restrict__ out_idx,
                                                             float *__restrict__ out_prob) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t g = draw_offset + (uint64_t)i;
    uint32_t c0 = (uint32_t)g, c1 = (uint32_t)(g >> 32), c2 = stream_id, c3 = 0u;
    philox4x32_10_st(c0, c1, c2, c3, (uint32_t)seed, (uint32_t)(seed >> 32));
    const uint64_t x = (uint64_t)c0 | ((uint64_t)c1 << 32);
    const double total = tree[1];
    double mass = __dmul_rn(__dmul_rn((double)(x >> 11), 0x1.0p-53), total);
    long long node = 1;
    while (node < P) {
        const double left = tree[2 * node];
        if (mass < left) {
            node = 2 * node;
        } else {
            mass = __dsub_rn(mass, left);
            node = 2 * node + 1;
        }
    }
    long long j = node - P;
    if (j >= size) j = size - 1;
    out_idx[i] = j;
    if (out_prob) out_prob[i] = (float)__ddiv_rn(tree[P + j], total);
}

}  // namespace b200rl

static bool is_pow2(int64_t x) { return x >= 1 && (x & (x - 1)) == 0; }

extern "C" int32_t b200rl_sumtree_update(double *tree, int64_t leaves, const int64_t *idx, const double *value,
                                         int64_t n, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(n >= 0, -2, "sumtree_update: n < 0");
    if (n == 0) return 0;
    B200RL_REQUIRE(tree && idx && value, -1, "sumtree_update: null pointer argument");
    B200RL_REQUIRE(is_pow2(leaves), -2, "sumtree_update: leaves=%lld is not a power of two", (long long)leaves);
    cudaStream_t s = (cudaStream_t)stream;
    B200RL_REQUIRE(leaves <= (1ll << 30), -3, "sumtree_update: leaves=%lld > 2^30", (long long)leaves);
    B200RL_REQUIRE(reinterpret_cast<uintptr_t>(tree) % 16 == 0, -4, "sumtree_update: tree must be 16-byte aligned");
    if (n <= kUpdMax) {
        int G = 1;
        while (G < kMaxSubtrees && leaves / (2 * G) >= kSub) G *= 2;  // subtrees of at least kSub leaves
        sumtree_update_kernel<<<G, kUpdThreads, 0, s>>>(tree, leaves, G, reinterpret_cast<const long long *>(idx), value, (int)n);
        if (G > 1) sumtree_top_kernel<<<1, kUpdThreads, 0, s>>>(tree, G);
        return check_launch("sumtree_update_kernel");
    }
    // bulk load: set the leaves, then rebuild every level
    sumtree_set_leaves_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(tree, leaves, reinterpret_cast<const long long *>(idx), value, n);
    for (int64_t width = leaves >> 1; width >= 1; width >>= 1)
        sumtree_level_kernel<<<(unsigned)((width + 255) / 256), 256, 0, s>>>(tree, width);
    return check_launch("sumtree_level_kernel");
}

extern "C" int32_t b200rl_sumtree_sample(const double *tree, int64_t leaves, int64_t size, uint64_t seed,
                                         uint32_t stream_id, uint64_t draw_offset, int64_t n, int64_t *out_idx,
                                         float *out_prob, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(n >= 0, -2, "sumtree_sample: n < 0");
    if (n == 0) return 0;
    B200RL_REQUIRE(tree && out_idx, -1, "sumtree_sample: null pointer argument");
    B200RL_REQUIRE(is_pow2(leaves), -2, "sumtree_sample: leaves=%lld is not a power of two", (long long)leaves);
    B200RL_REQUIRE(size >= 1 && size <= leaves, -2, "sumtree_sample: size=%lld outside [1, leaves]", (long long)size);
    sumtree_sample_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        tree, leaves, size, seed, stream_id, draw_offset, n, reinterpret_cast<long long *>(out_idx), out_prob);
    return check_launch("sumtree_sample_kernel");
}
