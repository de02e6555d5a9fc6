// Error slot, launch counter and version of the C-ABI (include/dctr.h).
#include <atomic>
#include <cstdarg>
#include <cstdlib>
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

// launch mode: the DCTR_PDL environment variable (0 / 1 / 2) wins; otherwise what the caller last set for this thread
// with dctr_set_launch_mode (the step engines choose per model and batch size); plain launches by default
static thread_local int g_pdl = 0;
int pdl_mode() {
    static const int env = [] { const char* e = getenv("DCTR_PDL"); return (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : -1; }();
    return env >= 0 ? env : g_pdl;
}

static thread_local int32_t* g_id_status = nullptr;
int32_t* id_status() { return g_id_status; }

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
int dctr_set_launch_mode(int mode) {
    if (mode < 0 || mode > 2) { dctr::set_error("set_launch_mode: mode %d not in 0..2", mode); return DCTR_EINVAL; }
    dctr::g_pdl = mode;
    return DCTR_OK;
}
int dctr_set_id_status(int32_t* status) { dctr::g_id_status = status; return DCTR_OK; }
const char* dctr_last_error(void) { return dctr::g_err; }
long long dctr_launch_count(void) { return dctr::g_launches.load(); }
}
