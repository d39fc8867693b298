// Peer-memory plumbing for K8 (one process per GPU): export a device buffer as a CUDA IPC handle, map a
// peer's buffer into this process.  The buffers stay owned by the caller's allocator (PyTorch); the handle
// names the cudaMalloc block the buffer lives in, so the byte offset inside that block travels with it.
// Mapping a peer's block enables peer access lazily (NVLink / NVSwitch on a B200 box).
#include <string.h>
#include "common.cuh"

namespace b200rl {

static CUresult address_range(CUdeviceptr *base, size_t *size, CUdeviceptr ptr) {
    typedef CUresult (*fn_t)(CUdeviceptr *, size_t *, CUdeviceptr);
    static fn_t fn = []() -> fn_t {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            return nullptr;
        }
        return reinterpret_cast<fn_t>(f);
    }();
    return fn ? fn(base, size, ptr) : CUDA_ERROR_NOT_SUPPORTED;
}

}  // namespace b200rl

extern "C" int32_t b200rl_ipc_export(const void *ptr, void *handle64, int64_t *offset) {
    using namespace b200rl;
    B200RL_REQUIRE(ptr && handle64 && offset, -1, "ipc_export: null pointer argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size is part of the ABI");
    CUdeviceptr base = 0;
    size_t size = 0;
    const CUresult r = address_range(&base, &size, reinterpret_cast<CUdeviceptr>(ptr));
    B200RL_REQUIRE(r == CUDA_SUCCESS, -5, "ipc_export: cuMemGetAddressRange failed (%d)", (int)r);
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, reinterpret_cast<void *>(base));
    if (e != cudaSuccess) {
        set_error("ipc_export: cudaIpcGetMemHandle: %s (buffers from an expandable-segments allocator cannot be shared)",
                  cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    memcpy(handle64, &h, 64);
    *offset = (int64_t)(reinterpret_cast<CUdeviceptr>(ptr) - base);
    return 0;
}

extern "C" int32_t b200rl_ipc_import(const void *handle64, int64_t offset, void **out_ptr) {
    using namespace b200rl;
    B200RL_REQUIRE(handle64 && out_ptr && offset >= 0, -1, "ipc_import: bad argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void *base = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        set_error("ipc_import: cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    *out_ptr = static_cast<char *>(base) + offset;
    return 0;
}

extern "C" int32_t b200rl_ipc_close(void *ptr, int64_t offset) {
    using namespace b200rl;
    B200RL_REQUIRE(ptr && offset >= 0, -1, "ipc_close: bad argument");
    const cudaError_t e = cudaIpcCloseMemHandle(static_cast<char *>(ptr) - offset);
    if (e != cudaSuccess) {
        set_error("ipc_close: cudaIpcCloseMemHandle: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return (int32_t)e;
    }
    return 0;
}
