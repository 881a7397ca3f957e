// Counter-based dropout masks of the training step (dropout.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

struct CssDrop {
    uint32_t key;     // mix(seed, site)
    uint32_t thr;     // floor(p * 2^24): an element is kept when its 24-bit hash is >= thr
    float scale;      // 1 / (1 - p)
};

__host__ __device__ __forceinline__ uint32_t css_dropout_mix(uint32_t x) {      // "lowbias32"
    x ^= x >> 16; x *= 0x7feb352dU;
    x ^= x >> 15; x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}
__device__ __forceinline__ bool css_dropout_keep(const CssDrop& d, uint32_t index) {
    return (css_dropout_mix(index ^ d.key) >> 8) >= d.thr;
}
uint32_t css_dropout_key(uint32_t seed, uint32_t site);
inline CssDrop css_make_drop(float p, uint32_t seed, uint32_t site) {
    CssDrop d = {css_dropout_key(seed, site), (uint32_t)((double)p * 16777216.0), 1.f / (1.f - p)};
    return d;
}
// site ids: 0 = embedding; block l: 1 + 8*l + k, k = 0 FFN-in hidden, 1 FFN-in output, 2 attention probabilities,
// 3 attention output, 4 conv module, 5 FFN-out hidden, 6 FFN-out output
inline uint32_t css_drop_site(int block, int k) { return block < 0 ? 0u : 1u + 8u * (uint32_t)block + (uint32_t)k; }

// dst = residual + alpha * act(drop(src)) over n contiguous floats (n % 4 == 0); dst may alias src / residual
int dropout_f32_launch(float* dst, const float* src, const float* residual, long n, const CssDrop& dr, float alpha, int relu,
                       cudaStream_t st);
// bf16 split planes in place; element index = row * cols + col
int dropout_planes_launch(void* planes, long plane_stride, int n_planes, long rows, int cols, int ld, const CssDrop& dr,
                          cudaStream_t st);
