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

// ---------------------------------------------------------------------------------------------
// Optional per-category kernel timing with CUDA events on the launching stream (bench.py's
// roofline numbers).  Off by default; when on, every instrumented launch is bracketed by two events.
// ---------------------------------------------------------------------------------------------
#include <vector>
struct ProfRec { cudaEvent_t a, b; int cat; double work; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;
static std::vector<ProfRec> g_prof_pool;

void css_prof_begin(int cat, double work, cudaStream_t st) {
    if (!g_prof_on) return;
    ProfRec r;
    if (!g_prof_pool.empty()) { r = g_prof_pool.back(); g_prof_pool.pop_back(); }
    else { cudaEventCreate(&r.a); cudaEventCreate(&r.b); }
    r.cat = cat; r.work = work;
    cudaEventRecord(r.a, st);
    g_prof.push_back(r);
}
void css_prof_end(cudaStream_t st) {
    if (!g_prof_on || g_prof.empty()) return;
    cudaEventRecord(g_prof.back().b, st);
}
extern "C" void css_profile_enable(int on) { g_prof_on = on != 0; }
// Synchronises the recorded events and accumulates per category: ms[cat], work[cat], count[cat].
extern "C" int css_profile_collect(double* ms, double* work, long* count, int n_cat) {
    for (int i = 0; i < n_cat; ++i) { ms[i] = 0; work[i] = 0; count[i] = 0; }
    for (ProfRec& r : g_prof) {
        float t = 0.f;
        if (cudaEventSynchronize(r.b) != cudaSuccess || cudaEventElapsedTime(&t, r.a, r.b) != cudaSuccess) {
            css_set_error("css_profile_collect: %s", cudaGetErrorString(cudaGetLastError()));
            return -3;
        }
        if (r.cat >= 0 && r.cat < n_cat) { ms[r.cat] += t; work[r.cat] += r.work; count[r.cat] += 1; }
        g_prof_pool.push_back(r);
    }
    g_prof.clear();
    return 0;
}
