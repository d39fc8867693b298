// K4/K5: device-resident replay -- uniform index sampling, row gather, ring append.
// Replaces rlmeta ReplayBuffer(CircularBuffer, UniformSampler) as used at reference
// agents/impala/builder.py:30-36 and agents/impala/learning.py:121-123,142 (sample ->
// RPC -> 5*B host-to-device copies -> torch.stack): storage lives in HBM, indices are
// drawn on the device and rows are copied HBM -> HBM straight into the [B,...] batch.
//
// Sampler stream (specified in oracle/replay_ref.py, pinned by the Random123 KATs):
// Philox4x32-10, ctr = (lo32(g), hi32(g), stream_id, 0), key = (lo32(seed), hi32(seed)),
// g = draw_offset + i;  index = mulhi64(out0 | out1 << 32, size).  Integer only: bit exact.
#include "common.cuh"

namespace b200rl {

__device__ __forceinline__ void philox4x32_10(uint32_t &c0, uint32_t &c1, uint32_t &c2, uint32_t &c3,
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

// draw number g of stream (seed, stream_id), reduced to [0, size)
__device__ __forceinline__ long long uniform_index(uint64_t seed, uint32_t stream_id, uint64_t g, uint64_t size) {
    uint32_t c0 = (uint32_t)g, c1 = (uint32_t)(g >> 32), c2 = stream_id, c3 = 0u;
    philox4x32_10(c0, c1, c2, c3, (uint32_t)seed, (uint32_t)(seed >> 32));
    const uint64_t x = (uint64_t)c0 | ((uint64_t)c1 << 32);
    return (long long)__umul64hi(x, size);
}

__global__ void __launch_bounds__(256) sample_uniform_kernel(uint64_t seed, uint32_t stream_id,
                                                             uint64_t draw_offset, long long n,
                                                             uint64_t size, long long *__restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = uniform_index(seed, stream_id, draw_offset + (uint64_t)i, size);
}

// Row copy with a per-row source or destination index.  The work is flattened over
// (row, vector) so that consecutive threads move consecutive 16-byte vectors of one row
// (fully coalesced on both sides) whatever the row size; each thread keeps kUnroll
// independent vectors in flight.
//   GATHER : dst[i]                    = src[clamp(idx[i])]
//   RING   : dst[(count + i) % cap]    = src[i]
enum { COPY_GATHER = 0, COPY_RING = 1, COPY_SAMPLE_GATHER = 2 };
constexpr int kCopyThreads = 256, kCopyUnroll = 4;

template <typename V, int MODE>
__global__ void __launch_bounds__(kCopyThreads) row_copy_kernel(
    const V *__restrict__ src, V *__restrict__ dst, const long long *__restrict__ idx, long long n,
    long long vec_per_row, long long nrows, long long count) {
    const long long total = n * vec_per_row;
    const long long base = ((long long)blockIdx.x * kCopyUnroll) * kCopyThreads + threadIdx.x;
    V regs[kCopyUnroll];
    long long d_off[kCopyUnroll];
#pragma unroll
    for (int u = 0; u < kCopyUnroll; ++u) {
        const long long k = base + (long long)u * kCopyThreads;
        d_off[u] = -1;
        if (k < total) {
            const long long row = k / vec_per_row, col = k - row * vec_per_row;
            long long s_row, d_row;
            if (MODE == COPY_GATHER) {
                long long j = __ldg(idx + row);
                j = j < 0 ? 0 : (j >= nrows ? nrows - 1 : j);
                s_row = j;
                d_row = row;
            } else {
                s_row = row;
                d_row = (count + row) % nrows;
            }
            regs[u] = __ldcs(src + s_row * vec_per_row + col);  // streaming: touched once
            d_off[u] = d_row * vec_per_row + col;
        }
    }
#pragma unroll
    for (int u = 0; u < kCopyUnroll; ++u)
        if (d_off[u] >= 0) __stcs(dst + d_off[u], regs[u]);
}

template <int MODE>
static int32_t launch_row_copy(const void *src, void *dst, const long long *idx, long long n,
                               long long row_bytes, long long nrows, long long count, cudaStream_t stream) {
    if (n == 0 || row_bytes == 0) return 0;
    const uintptr_t align = reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst) | (uintptr_t)row_bytes;
    const int vec = (align % 16 == 0) ? 16 : (align % 8 == 0) ? 8 : (align % 4 == 0) ? 4 : 1;
    const long long vpr = row_bytes / vec, total = n * vpr;
    const long long per_cta = (long long)kCopyThreads * kCopyUnroll;
    const long long grid = (total + per_cta - 1) / per_cta;
    B200RL_REQUIRE(grid < (1ll << 31), -3, "row copy: too many elements");
    switch (vec) {
        case 16: row_copy_kernel<int4, MODE><<<(unsigned)grid, kCopyThreads, 0, stream>>>((const int4 *)src, (int4 *)dst, idx, n, vpr, nrows, count); break;
        case 8: row_copy_kernel<int2, MODE><<<(unsigned)grid, kCopyThreads, 0, stream>>>((const int2 *)src, (int2 *)dst, idx, n, vpr, nrows, count); break;
        case 4: row_copy_kernel<int, MODE><<<(unsigned)grid, kCopyThreads, 0, stream>>>((const int *)src, (int *)dst, idx, n, vpr, nrows, count); break;
        default: row_copy_kernel<char, MODE><<<(unsigned)grid, kCopyThreads, 0, stream>>>((const char *)src, (char *)dst, idx, n, vpr, nrows, count); break;
    }
    return check_launch("row_copy_kernel");
}

}  // namespace b200rl

extern "C" int32_t b200rl_sample_uniform(uint64_t seed, uint32_t stream_id, uint64_t draw_offset, int64_t n,
                                         int64_t size, int64_t *out_idx, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(n >= 0, -2, "sample_uniform: n < 0");
    if (n == 0) return 0;
    B200RL_REQUIRE(out_idx, -1, "sample_uniform: null output");
    B200RL_REQUIRE(size >= 1, -2, "sample_uniform: cannot sample from an empty buffer (size=%lld)", (long long)size);
    sample_uniform_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        seed, stream_id, draw_offset, n, (uint64_t)size, reinterpret_cast<long long *>(out_idx));
    return check_launch("sample_uniform_kernel");
}

extern "C" int32_t b200rl_gather_rows(const void *src, int64_t row_bytes, const int64_t *idx, int64_t n,
                                      int64_t nrows, void *dst, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(n >= 0 && row_bytes >= 0, -2, "gather_rows: negative size");
    if (n == 0 || row_bytes == 0) return 0;
    B200RL_REQUIRE(src && idx && dst, -1, "gather_rows: null pointer argument");
    B200RL_REQUIRE(nrows >= 1, -2, "gather_rows: empty source");
    return launch_row_copy<COPY_GATHER>(src, dst, reinterpret_cast<const long long *>(idx), n, row_bytes, nrows, 0,
                                        (cudaStream_t)stream);
}

extern "C" int32_t b200rl_ring_append(void *storage, int64_t row_bytes, int64_t capacity, int64_t count,
                                      const void *rows, int64_t n, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(n >= 0 && row_bytes >= 0 && count >= 0, -2, "ring_append: negative size");
    if (n == 0 || row_bytes == 0) return 0;
    B200RL_REQUIRE(storage && rows, -1, "ring_append: null pointer argument");
    B200RL_REQUIRE(capacity >= 1 && n <= capacity, -2, "ring_append: n=%lld exceeds capacity=%lld", (long long)n, (long long)capacity);
    return launch_row_copy<COPY_RING>(rows, storage, nullptr, n, row_bytes, capacity, count, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------
// N2: multi-field ring append in ONE launch, with the actor-side `discount = !done * gamma`
// (reference ImpalaActor._make_replay, agents/impala/learning.py:77-80) fused into the copy.
// Each field owns a contiguous range of CTAs sized for it and is flattened over (row, 16-byte
// vector) like row_copy_kernel.
namespace b200rl {

constexpr int kMaxFields = 8;
struct FieldTable {
    const void *src[kMaxFields];
    void *dst[kMaxFields];
    long long row_bytes[kMaxFields];  // bytes per SOURCE row
    int kind[kMaxFields];             // B200RL_FIELD_*
    int vec[kMaxFields];              // copy width in bytes (16/8/4/1); discount fields: 1
    float gamma[kMaxFields];
    long long n, capacity, count;
    const long long *idx;  // gather only
    // sample + gather in one launch (COPY_SAMPLE_GATHER): row i comes from storage row uniform_index(draw_offset + i)
    uint64_t seed, draw_offset, size;
    uint32_t stream_id;
    long long *idx_out;
    const unsigned long long *draw_counter;  // nullable, device: added to draw_offset (CUDA-graph replay, see b200rl.h)
    unsigned int cta_end[kMaxFields];  // field f owns CTAs [cta_end[f-1], cta_end[f])
    int n_fields;
};

// MODE COPY_RING: dst[(count + i) % capacity] = src[i];  COPY_GATHER: dst[i] = src[clamp(idx[i])];
// COPY_SAMPLE_GATHER: dst[i] = src[uniform_index(draw_offset + i)] and idx_out[i] = that index -- K4 and K5 in ONE launch:
// a CTA first draws the indices of the few rows its 1,024 vectors touch (3 for 12 KB rows) into shared memory, so the
// sampler's launch (3.4 us of the 19 us sample + gather step at BASELINE config 4) and the index round trip through HBM
// disappear; the draws are the same Philox stream, bit for bit.
template <int MODE>
__global__ void __launch_bounds__(kCopyThreads) fields_copy_kernel(const FieldTable ft) {
    int f = 0;
    while (f < ft.n_fields - 1 && blockIdx.x >= ft.cta_end[f]) ++f;  // every field gets exactly the CTAs it needs
    const unsigned int cta = blockIdx.x - (f ? ft.cta_end[f - 1] : 0u);
    const long long rb = ft.row_bytes[f];
    const int vec = ft.vec[f];
    const long long vpr = rb / vec, total = ft.n * vpr;
    const long long base = ((long long)cta * kCopyUnroll) * kCopyThreads + threadIdx.x;
    __shared__ long long s_idx[MODE == COPY_SAMPLE_GATHER ? kCopyThreads * kCopyUnroll + 1 : 1];
    long long first_row = 0;
    if (MODE == COPY_SAMPLE_GATHER) {
        // launched with the PDL attribute: the draws are pure arithmetic and run while the preceding kernel drains;
        // nothing in global memory is touched before griddepcontrol.wait (storage may have just been appended to, and
        // the previous consumer of idx_out may still be reading it)
        pdl_launch_dependents();
        const long long k0 = (long long)cta * kCopyUnroll * kCopyThreads;                  // < total: the grid is exact
        const long long k1 = min(total, k0 + (long long)kCopyUnroll * kCopyThreads) - 1;
        first_row = k0 / vpr;
        const long long last_row = k1 / vpr;
        uint64_t off = ft.draw_offset;
        if (ft.draw_counter) {  // the counter is advanced by a preceding launch of this stream: read it after the wait
            pdl_wait();
            off += *reinterpret_cast<const volatile unsigned long long *>(ft.draw_counter);
        }
        for (long long r = first_row + threadIdx.x; r <= last_row; r += kCopyThreads)
            s_idx[r - first_row] = uniform_index(ft.seed, ft.stream_id, off + (uint64_t)r, ft.size);
        __syncthreads();
        if (!ft.draw_counter) {
            // the source addresses are known before the dependency wait: ask L2 for this CTA's part of its rows
            // (cp.async.bulk.prefetch.L2 -- a hint, never a read; 16-byte vectors only: the bulk unit's alignment)
            if (vec == 16 && threadIdx.x <= last_row - first_row) {
                const long long r = first_row + threadIdx.x;
                const long long ka = max(k0, r * vpr), kb = min(k1 + 1, (r + 1) * vpr);
                long long j = s_idx[threadIdx.x];
                j = j < 0 ? 0 : (j >= ft.capacity ? ft.capacity - 1 : j);
                bulk_prefetch_l2(static_cast<const char *>(ft.src[f]) + (j * vpr + (ka - r * vpr)) * 16, (uint32_t)((kb - ka) * 16));
            }
            pdl_wait();
        }
        if (f == 0)
            for (long long r = first_row + threadIdx.x; r <= last_row; r += kCopyThreads)
                if (r * vpr >= k0) ft.idx_out[r] = s_idx[r - first_row];  // every row starts in exactly one CTA of field 0
    }
    if (base >= total) return;
    if (MODE == COPY_RING && ft.kind[f] == B200RL_FIELD_DONE_TO_DISCOUNT) {
        const unsigned char *src = static_cast<const unsigned char *>(ft.src[f]);
        float *dst = static_cast<float *>(ft.dst[f]);
        const float g = ft.gamma[f];
#pragma unroll
        for (int u = 0; u < kCopyUnroll; ++u) {
            const long long k = base + (long long)u * kCopyThreads;
            if (k < total) {
                const long long row = k / rb, col = k - row * rb;
                const long long d_row = (ft.count + row) % ft.capacity;
                dst[d_row * rb + col] = src[k] ? 0.f : g;  // logical_not(done) * gamma
            }
        }
        return;
    }
    auto copy = [&](auto tag) {
        using V = decltype(tag);
        const V *src = static_cast<const V *>(ft.src[f]);
        V *dst = static_cast<V *>(ft.dst[f]);
        V regs[kCopyUnroll];
        long long d_off[kCopyUnroll];
#pragma unroll
        for (int u = 0; u < kCopyUnroll; ++u) {
            const long long k = base + (long long)u * kCopyThreads;
            d_off[u] = -1;
            if (k < total) {
                const long long row = k / vpr, col = k - row * vpr;
                if (MODE == COPY_RING) {
                    regs[u] = __ldcs(src + k);
                    d_off[u] = ((ft.count + row) % ft.capacity) * vpr + col;
                } else {
                    long long j = MODE == COPY_SAMPLE_GATHER ? s_idx[row - first_row] : __ldg(ft.idx + row);
                    j = j < 0 ? 0 : (j >= ft.capacity ? ft.capacity - 1 : j);
                    regs[u] = __ldcs(src + j * vpr + col);
                    d_off[u] = k;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < kCopyUnroll; ++u)
            if (d_off[u] >= 0) __stcs(dst + d_off[u], regs[u]);
    };
    switch (vec) {
        case 16: copy(int4{}); break;
        case 8: copy(int2{}); break;
        case 4: copy(int{}); break;
        default: copy(char{}); break;
    }
}

}  // namespace b200rl

static int32_t fill_field_table(const char *who, const b200rl_field_t *fields, int32_t n_fields, long long n,
                                bool allow_discount, b200rl::FieldTable &ft, long long &max_ctas) {
    using namespace b200rl;
    max_ctas = 0;  // total CTAs over all fields
    ft.n_fields = n_fields;
    for (int f = 0; f < n_fields; ++f) {
        const b200rl_field_t &x = fields[f];
        B200RL_REQUIRE(x.src && x.dst, -1, "%s: field %d has a null pointer", who, f);
        B200RL_REQUIRE(x.row_bytes >= 1, -2, "%s: field %d row_bytes=%lld", who, f, (long long)x.row_bytes);
        B200RL_REQUIRE(x.kind == B200RL_FIELD_COPY || (allow_discount && x.kind == B200RL_FIELD_DONE_TO_DISCOUNT), -2,
                       "%s: field %d has unsupported kind %d", who, f, x.kind);
        ft.src[f] = x.src; ft.dst[f] = x.dst; ft.row_bytes[f] = x.row_bytes; ft.kind[f] = x.kind; ft.gamma[f] = x.gamma;
        int vec = 1;
        if (x.kind == B200RL_FIELD_COPY) {
            const uintptr_t al = reinterpret_cast<uintptr_t>(x.src) | reinterpret_cast<uintptr_t>(x.dst) | (uintptr_t)x.row_bytes;
            vec = (al % 16 == 0) ? 16 : (al % 8 == 0) ? 8 : (al % 4 == 0) ? 4 : 1;
        }
        ft.vec[f] = vec;
        const long long total = n * (x.row_bytes / vec), per_cta = (long long)kCopyThreads * kCopyUnroll;
        max_ctas += (total + per_cta - 1) / per_cta;
        ft.cta_end[f] = (unsigned int)max_ctas;
    }
    B200RL_REQUIRE(max_ctas < (1ll << 31), -3, "%s: too many elements", who);
    return 0;
}

extern "C" int32_t b200rl_ring_append_fields(const b200rl_field_t *fields, int32_t n_fields, int64_t capacity,
                                             int64_t count, int64_t n, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(n >= 0 && count >= 0, -2, "ring_append_fields: negative size");
    B200RL_REQUIRE(n_fields >= 1 && n_fields <= kMaxFields, -2, "ring_append_fields: 1..%d fields (got %d)", kMaxFields, n_fields);
    B200RL_REQUIRE(fields, -1, "ring_append_fields: null field table");
    if (n == 0) return 0;
    B200RL_REQUIRE(capacity >= 1 && n <= capacity, -2, "ring_append_fields: n=%lld exceeds capacity=%lld", (long long)n, (long long)capacity);
    FieldTable ft = {};
    ft.n = n; ft.capacity = capacity; ft.count = count; ft.idx = nullptr;
    long long max_ctas;
    if (int32_t rc = fill_field_table("ring_append_fields", fields, n_fields, n, true, ft, max_ctas)) return rc;
    fields_copy_kernel<COPY_RING><<<(unsigned)max_ctas, kCopyThreads, 0, (cudaStream_t)stream>>>(ft);
    return check_launch("fields_copy_kernel<ring>");
}

extern "C" int32_t b200rl_gather_fields(const b200rl_field_t *fields, int32_t n_fields, const int64_t *idx,
                                        int64_t n, int64_t nrows, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(n >= 0, -2, "gather_fields: negative size");
    B200RL_REQUIRE(n_fields >= 1 && n_fields <= kMaxFields, -2, "gather_fields: 1..%d fields (got %d)", kMaxFields, n_fields);
    B200RL_REQUIRE(fields, -1, "gather_fields: null field table");
    if (n == 0) return 0;
    B200RL_REQUIRE(idx, -1, "gather_fields: null index");
    B200RL_REQUIRE(nrows >= 1, -2, "gather_fields: empty source");
    FieldTable ft = {};
    ft.n = n; ft.capacity = nrows; ft.count = 0; ft.idx = reinterpret_cast<const long long *>(idx);
    long long max_ctas;
    if (int32_t rc = fill_field_table("gather_fields", fields, n_fields, n, false, ft, max_ctas)) return rc;
    fields_copy_kernel<COPY_GATHER><<<(unsigned)max_ctas, kCopyThreads, 0, (cudaStream_t)stream>>>(ft);
    return check_launch("fields_copy_kernel<gather>");
}

// K4 + K5 in one launch (the learner's ReplayBuffer.sample with the uniform sampler, reference
// agents/impala/learning.py:121 / agents/ppo/learning.py:112): draws n indices in [0, size) exactly as
// b200rl_sample_uniform(seed, stream_id, draw_offset, n, size) does, writes them to out_idx and gathers those rows of every field.
extern "C" int32_t b200rl_sample_gather_fields(const b200rl_field_t *fields, int32_t n_fields, uint64_t seed,
                                               uint32_t stream_id, uint64_t draw_offset, const uint64_t *draw_counter,
                                               int64_t n, int64_t size, int64_t nrows, int64_t *out_idx, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(n >= 0, -2, "sample_gather_fields: negative size");
    B200RL_REQUIRE(n_fields >= 1 && n_fields <= kMaxFields, -2, "sample_gather_fields: 1..%d fields (got %d)", kMaxFields, n_fields);
    B200RL_REQUIRE(fields, -1, "sample_gather_fields: null field table");
    if (n == 0) return 0;
    B200RL_REQUIRE(out_idx, -1, "sample_gather_fields: null index output");
    B200RL_REQUIRE(size >= 1 && size <= nrows, -2,
                   "sample_gather_fields: cannot sample %lld of %lld rows (size must be in 1..nrows)", (long long)size, (long long)nrows);
    FieldTable ft = {};
    ft.n = n; ft.capacity = nrows; ft.count = 0; ft.idx = nullptr;
    ft.seed = seed; ft.stream_id = stream_id; ft.draw_offset = draw_offset; ft.size = (uint64_t)size;
    ft.idx_out = reinterpret_cast<long long *>(out_idx);
    ft.draw_counter = reinterpret_cast<const unsigned long long *>(draw_counter);
    long long max_ctas;
    if (int32_t rc = fill_field_table("sample_gather_fields", fields, n_fields, n, false, ft, max_ctas)) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)max_ctas);
    cfg.blockDim = dim3(kCopyThreads);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // PDL: see pdl_wait() in the kernel
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, fields_copy_kernel<COPY_SAMPLE_GATHER>, ft);
    if (e != cudaSuccess) {
        set_error("fields_copy_kernel<sample_gather> launch: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    return check_launch("fields_copy_kernel<sample_gather>");
}

namespace b200rl {
__global__ void counter_add_kernel(unsigned long long *counter, long long delta) { *counter += (unsigned long long)delta; }
}  // namespace b200rl

// *counter += delta in stream order (one thread).  The device-resident draw counter of a sampler whose launches are replayed
// from a CUDA graph: the sampling launch reads it, this launch advances it for the next replay.
extern "C" int32_t b200rl_counter_add(uint64_t *counter, int64_t delta, void *stream) {
    using namespace b200rl;
    B200RL_REQUIRE(counter, -1, "counter_add: null counter");
    counter_add_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(reinterpret_cast<unsigned long long *>(counter), (long long)delta);
    return check_launch("counter_add_kernel");
}
