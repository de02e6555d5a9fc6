// Error slot, launch counter and version of the C-ABI (include/dctr.h).
#include <atomic>
#include <cstdarg>
#include "common.cuh"

namespace dctr {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// Called right after a <<<>>> launch: picks up launch-configuration errors without synchronising.
int check_launch(const char* what) {
    count_launch(1);
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("%s: %s", what, cudaGetErrorString(e));
        return DCTR_ECUDA;
    }
    return DCTR_OK;
}

}  // namespace dctr

extern "C" {
int dctr_version(void) { return DCTR_VERSION; }
const char* dctr_last_error(void) { return dctr::g_err; }
long long dctr_launch_count(void) { return dctr::g_launches.load(); }
}
