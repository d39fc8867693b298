// Error plumbing and version of the b200rl C ABI.
#include <stdarg.h>
#include <stdlib.h>
#include "common.cuh"

namespace b200rl {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int32_t check_launch(const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return (int32_t)e;
    }
    return 0;
}

// 2-D fp32 tensor map: `rows` rows of `inner` elements, row pitch `pitch_bytes`, box = box_rows x box_inner.
// Returns false (and leaves *map untouched) when the driver entry point is missing or the shape violates a
// TMA constraint (16-byte base/pitch/box alignment, box dimensions <= 256); callers then use the 1-D path.
bool encode_tmap_2d_f32(CUtensorMap *map, const void *base, uint64_t inner, uint64_t rows, uint64_t pitch_bytes,
                        uint32_t box_inner, uint32_t box_rows) {
    typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn fn = []() -> encode_fn {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            return nullptr;
        }
        return reinterpret_cast<encode_fn>(f);
    }();
    if (!fn) return false;
    if (reinterpret_cast<uintptr_t>(base) % 16 || pitch_bytes % 16 || (box_inner * 4u) % 16 || box_inner > 256 ||
        box_rows > 256 || box_inner == 0 || box_rows == 0 || inner == 0 || rows == 0)
        return false;
    const cuuint64_t dims[2] = {inner, rows};
    const cuuint64_t strides[1] = {pitch_bytes};
    const cuuint32_t box[2] = {box_inner, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void *>(base), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

}  // namespace b200rl

extern "C" int32_t b200rl_version(void) { return B200RL_VERSION; }
extern "C" const char *b200rl_last_error(void) { return b200rl::g_err; }
extern "C" int64_t b200rl_loss_workspace_bytes(void) { return b200rl::kWorkspaceBytes; }
extern "C" int64_t b200rl_trace_offset_bytes(void) { return b200rl::kTraceOffset; }
