// b200rl_vtrace_impala_loss_fwdbwd_noalloc: K1 for callers that cannot allocate (code under CUDA-graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  NO counterpart in the reference.
// Same kernels and arithmetic as b200rl_vtrace_impala_loss_fwdbwd; the only difference is where the scalar
// reduction's scratch (the "loss workspace": launch generation + one tagged word per CTA and scalar) lives -- here in a
// file-scope device array instead of a caller-allocated buffer.
//
//   * g_noalloc_scratch: kNoAllocScratchBytes = 256 + 16384 * 8 * 8 = 1,048,832 bytes (1 MiB + 256 B), held by EVERY GPU
//     that loads the library until the process exits.  That is the layout the reduction addresses (header, then
//     [8 scalars][16384 CTAs] words); the trace area of the ordinary workspace is NOT part of it, so B200RL_TRACE is refused.
//   * Limits, checked on the host before anything is launched: B <= 16384 batch columns per call (the tile kernel then
//     never needs more than 16384 CTAs' words), A <= 32, T as the ordinary entry point.
//   * In HOST code the array's name is a host-side stand-in, not a device address.  The device address is taken once per
//     device, with that device current, by cudaGetSymbolAddress (checked), cached, and passed to the kernel as its workspace.
//   * Every call on a device SHARES the array and finds the last call's contents in it.  The launch generation must
//     survive from call to call (the tags of a new launch are generation + 1, so old words can never be mistaken for new
//     ones); the "poisoned" word must start at zero for every call -- a timed-out earlier caller must not poison a resource
//     nobody can re-zero -- and is cleared by a one-thread kernel ahead of the loss kernel, in stream order.
//   * Calls that OVERLAP on the GPU (two streams, two host threads, a graph replayed while another call runs) RACE on
//     the array: scalars[] of both are undefined (gradients are unaffected).  Callers that need overlap use the ordinary
//     entry point with one workspace per stream.
//   * Nothing here calls cudaMalloc / cudaFree or an allocating library routine.
#include <mutex>
#include "impala_loss.cuh"

namespace b200rl {

constexpr int64_t kNoAllocScratchBytes = kTraceOffset;  // header + partial words; see above
constexpr int64_t kNoAllocMaxColumns = 16384;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// Device code may name the array directly.  Clears what must start at zero for a call: the poisoned word (uint32 #1 of
// the header).  The generation (uint32 #0) and the tagged words are deliberately left alone.
__global__ void noalloc_begin_call_kernel() {
    reinterpret_cast<unsigned int *>(g_noalloc_scratch)[1] = 0u;
}

// Device address of the scratch on the CURRENT device (nullptr + error text on failure).
static void *noalloc_scratch_address() {
    static void *addr[kMaxDevices];
    static std::mutex mu;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        (void)cudaGetLastError();
        set_error("impala_loss_noalloc: cudaGetDevice failed or device index %d beyond %d", dev, kMaxDevices);
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);  // device `dev` is current
        if (e != cudaSuccess || !p) {
            (void)cudaGetLastError();
            set_error("impala_loss_noalloc: cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
            return nullptr;
        }
        addr[dev] = p;
    }
    return addr[dev];
}

}  // namespace b200rl

extern "C" int32_t b200rl_vtrace_impala_loss_fwdbwd_noalloc(
    const float *logits, const float *pi_ref, const int64_t *actions, const float *rewards,
    const float *discounts, const float *values, int64_t B, int64_t T, int64_t A, int64_t stride_b,
    int64_t stride_t, int64_t n_global_b, double lambda_, float clip_rho, float clip_pg_rho,
    float entropy_coeff, uint32_t flags, float *dlogits, float *dvalues, float *scalars, void *stream) {
    using namespace b200rl;
    static_assert(kNoAllocScratchBytes == B200RL_NOALLOC_SCRATCH_BYTES && kNoAllocMaxColumns == B200RL_NOALLOC_MAX_COLUMNS,
                  "include/b200rl.h states the limits");
    B200RL_REQUIRE(logits && pi_ref && actions && rewards && discounts && values && dlogits && dvalues && scalars, -1,
                   "impala_loss_noalloc: null pointer argument");
    B200RL_REQUIRE(B >= 1 && T >= 2 && A >= 1, -2, "impala_loss_noalloc: need B>=1, T>=2, A>=1 (got B=%lld T=%lld A=%lld)",
                   (long long)B, (long long)T, (long long)A);
    B200RL_REQUIRE(B <= kNoAllocMaxColumns, -3,
                   "impala_loss_noalloc: B=%lld batch columns; the built-in scratch (%lld bytes) serves at most %lld per call -- "
                   "use b200rl_vtrace_impala_loss_fwdbwd with a caller-allocated workspace for larger batches",
                   (long long)B, (long long)kNoAllocScratchBytes, (long long)kNoAllocMaxColumns);
    B200RL_REQUIRE(!(flags & B200RL_TRACE), -3,
                   "impala_loss_noalloc: B200RL_TRACE needs the trace area of a caller-allocated workspace");
    void *scratch = noalloc_scratch_address();
    if (!scratch) return (int32_t)cudaErrorInvalidSymbol;
    noalloc_begin_call_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>();
    if (int32_t rc = check_launch("noalloc_begin_call_kernel")) return rc;
    return b200rl_vtrace_impala_loss_fwdbwd(logits, pi_ref, actions, rewards, discounts, values, B, T, A, stride_b, stride_t,
                                            n_global_b, lambda_, clip_rho, clip_pg_rho, entropy_coeff, flags, dlogits,
                                            dvalues, scalars, scratch, stream);
}
