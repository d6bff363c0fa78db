// Error plumbing and version for libdlrmx (C ABI declared in include/dlrmx.h).
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace dlrmx {
static thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};

static std::atomic<int> g_tunables[DLRMX_TUNE_COUNT];
int get_tunable(int which) { return g_tunables[which].load(std::memory_order_relaxed); }
void set_tunable(int which, int v) { g_tunables[which].store(v); }

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace dlrmx

extern "C" {
int dlrmx_version(void) { return DLRMX_VERSION; }
const char* dlrmx_last_error(void) { return dlrmx::g_err; }
int64_t dlrmx_launch_count(void) { return dlrmx::g_launches.load(); }
int dlrmx_set_tunable(int which, int value) {
    if (which < 0 || which >= dlrmx::DLRMX_TUNE_COUNT) return DLRMX_E_ARG;
    dlrmx::set_tunable(which, value);
    return DLRMX_OK;
}
}
