// Error plumbing of the C-ABI: every entry point returns 0 or a negative code and leaves a
// message retrievable with css_last_error().  (The reference signals errors with Python
// exceptions -- SURVEY.md 8b "Errors"; the host wrappers turn non-zero codes into RuntimeError.)
#include "common.cuh"
#include <stdarg.h>
#include <string.h>

static thread_local char g_err[512] = "";
static unsigned long long g_launches = 0;   // kernels launched through the library

void css_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int css_check_launch(const char* what) {
    ++g_launches;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        css_set_error("%s: CUDA launch failed: %s", what, cudaGetErrorString(e));
        return -2;
    }
    return 0;
}

extern "C" const char* css_last_error(void) { return g_err; }

extern "C" int css_version(void) { return 100; }

// Number of kernels launched through the library since the last reset (bench.py gpu_launches).
extern "C" unsigned long long css_launch_count(int reset) {
    unsigned long long v = g_launches;
    if (reset) g_launches = 0;
    return v;
}
