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

int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

}  // namespace b200rl

extern "C" int32_t b200rl_version(void) { return B200RL_VERSION; }
extern "C" const char *b200rl_last_error(void) { return b200rl::g_err; }
extern "C" int64_t b200rl_loss_workspace_bytes(void) { return b200rl::kWorkspaceBytes; }
extern "C" int64_t b200rl_trace_offset_bytes(void) { return b200rl::kTraceOffset; }
