// runtime.cu — error text, launch counter, version (no kernels).
#include "common.cuh"
#include <atomic>
#include <cstdarg>
#include <cstring>

namespace lodgs {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
int check_cuda(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return LODGS_E_CUDA;
}

}  // namespace lodgs

extern "C" const char* lodgs_last_error(void) { return lodgs::g_err; }
extern "C" int lodgs_version(void) { return 100; }
extern "C" int lodgs_compiled_sm(void) { return 100; }
extern "C" uint64_t lodgs_launch_count(void) { return lodgs::g_launches.load(); }
