// Shared helpers for the css_b200 sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CSS_NBINS 257   // n_fft/2+1 for the 512-point STFT (conf/*.yaml frame_length 512)
#define CSS_KP 264      // padded bin pitch (elements) of every "bin-fastest" array
#define CSS_NFFT 512
#define CSS_HOP 256

// error plumbing (api.cu)
void css_set_error(const char* fmt, ...);
int css_check_launch(const char* what);
// per-category event timing (api.cu); categories: 0 gemm(flops) 1 attention(flops) 2 conv module(bytes)
// 3 layernorm(bytes) 4 lstm step(flops)
void css_prof_begin(int cat, double work, cudaStream_t st);
void css_prof_end(cudaStream_t st);

// categories 5.. are the HBM-bound DSP kernels, work = algorithmic bytes: 5 stft, 6 features + input MVN,
// 7 DOA merge, 8 stitch (pair decisions + connect), 9 MVDR (covariances, solve, apply, rescale), 10 iSTFT placement + peak
struct CssProfScope {
    cudaStream_t st;
    CssProfScope(int cat, double work, cudaStream_t s) : st(s) { css_prof_begin(cat, work, s); }
    ~CssProfScope() { css_prof_end(st); }
};

// Function attributes (the > 48 KB shared-memory opt-in) and SM counts are per DEVICE: caches are keyed by the current
// device ordinal so that a second GPU in the same process gets its own setup.
#define CSS_MAX_DEVICES 64
static inline int css_device() {
    int d = 0;
    cudaGetDevice(&d);
    return d & (CSS_MAX_DEVICES - 1);
}
struct CssPerDevice {
    bool done[CSS_MAX_DEVICES] = {};
    bool first() {
        const int d = css_device();
        if (done[d]) return false;
        done[d] = true;
        return true;
    }
};

#define CSS_REQUIRE(cond, ...)            \
    do {                                  \
        if (!(cond)) {                    \
            css_set_error(__VA_ARGS__);   \
            return -1;                    \
        }                                 \
    } while (0)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
// a * conj(b)
__device__ __forceinline__ float2 cmulc(float2 a, float2 b) {
    return make_float2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y);
}

static inline int ceil_div(long a, long b) { return (int)((a + b - 1) / b); }
