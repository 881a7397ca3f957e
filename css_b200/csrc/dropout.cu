// Dropout of the training step (css/models/conformer.py:127,249,254,302,317,319: nn.Dropout(p) at six sites per network /
// block), as counter-based masks: keep(element) is a pure function of (seed, site, element index), so the backward
// re-derives the mask of the forward instead of storing it.
//
//   key  = mix(seed * 0x9E3779B1 + site)                      (host)
//   keep = (mix(index ^ key) >> 8) >= floor(p * 2^24)         mix = the 32-bit "lowbias32" integer hash
//   y    = keep ? x / (1 - p) : 0                             (inverted dropout, as torch.nn.Dropout)
//
// The reference draws its masks from PyTorch's Philox stream, which cannot be reproduced outside PyTorch: parity for
// p > 0 is defined against the SAME masks (tests/helpers.py:dropout_mask mirrors this hash in numpy and feeds them to
// the oracle's autograd), and for p = 0 against the reference itself.
#include "common.cuh"
#include "dropout.h"
#include <cuda_bf16.h>

uint32_t css_dropout_key(uint32_t seed, uint32_t site) { return css_dropout_mix(seed * 0x9E3779B1u + site); }

// dst[i] = (residual ? residual[i] : 0) + alpha * act(drop(src[i]))   over n = rows*cols contiguous elements; relu optional.
// dst may alias src or residual.
__global__ void __launch_bounds__(256)
dropout_f32_kernel(float* dst, const float* src, const float* residual, long n4, CssDrop dr, float alpha, int relu) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
        const float4 v = reinterpret_cast<const float4*>(src)[i];
        float4 r = residual ? reinterpret_cast<const float4*>(residual)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        const uint32_t e = (uint32_t)(i * 4);
        float o[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            o[k] = css_dropout_keep(dr, e + k) ? o[k] * dr.scale : 0.f;
            if (relu) o[k] = fmaxf(o[k], 0.f);
        }
        r.x += alpha * o[0]; r.y += alpha * o[1]; r.z += alpha * o[2]; r.w += alpha * o[3];
        reinterpret_cast<float4*>(dst)[i] = r;
    }
}

int dropout_f32_launch(float* dst, const float* src, const float* residual, long n, const CssDrop& dr, float alpha, int relu,
                       cudaStream_t st) {
    CSS_REQUIRE(n > 0 && (n & 3) == 0 && n < (1L << 32), "dropout: element count %ld must be a multiple of 4 below 2^32", n);
    long g = (n / 4 + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    dropout_f32_kernel<<<(unsigned)g, 256, 0, st>>>(dst, src, residual, n / 4, dr, alpha, relu);
    return css_check_launch("dropout_f32");
}

// bf16 split planes (rows, ld) in place: v = (p0 + p1 + ..) -> drop(v) re-split.  Element index = row * cols + col.
__global__ void __launch_bounds__(256)
dropout_planes_kernel(__nv_bfloat16* planes, long plane_stride, int n_planes, long rows, int cols, int ld, CssDrop dr) {
    const long total = rows * (cols / 2);
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long r = i / (cols / 2);
        const int c = (int)(i - r * (cols / 2)) * 2;
        float a = 0.f, b = 0.f;
        for (int p = 0; p < n_planes; ++p) {
            const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(planes + p * plane_stride + r * ld + c);
            a += __bfloat162float(h.x);
            b += __bfloat162float(h.y);
        }
        const uint32_t e = (uint32_t)(r * cols + c);
        a = css_dropout_keep(dr, e) ? a * dr.scale : 0.f;
        b = css_dropout_keep(dr, e + 1) ? b * dr.scale : 0.f;
        for (int p = 0; p < n_planes; ++p) {
            __nv_bfloat162 h;
            h.x = __float2bfloat16_rn(a);
            h.y = __float2bfloat16_rn(b);
            a -= __bfloat162float(h.x);
            b -= __bfloat162float(h.y);
            *reinterpret_cast<__nv_bfloat162*>(planes + p * plane_stride + r * ld + c) = h;
        }
    }
}

int dropout_planes_launch(void* planes, long plane_stride, int n_planes, long rows, int cols, int ld, const CssDrop& dr,
                          cudaStream_t st) {
    CSS_REQUIRE(rows > 0 && cols > 0 && (cols & 1) == 0 && ld >= cols && rows * cols < (1L << 32), "dropout_planes: bad shape");
    long g = (rows * (cols / 2) + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    dropout_planes_kernel<<<(unsigned)g, 256, 0, st>>>(reinterpret_cast<__nv_bfloat16*>(planes), plane_stride, n_planes, rows, cols, ld, dr);
    return css_check_launch("dropout_planes");
}

// test / tooling entry: mask[i] = keep(i) for i < n (uint8), the mask the kernels above apply at (seed, site)
__global__ void dropout_mask_kernel(unsigned char* mask, long n, CssDrop dr) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
        mask[i] = css_dropout_keep(dr, (uint32_t)i) ? 1 : 0;
}
extern "C" int css_dropout_mask(unsigned char* mask, long n, float p, unsigned seed, unsigned site, cudaStream_t st) {
    CSS_REQUIRE(n > 0 && n < (1L << 32) && p >= 0.f && p < 1.f, "css_dropout_mask: bad arguments");
    long g = (n + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    dropout_mask_kernel<<<(unsigned)g, 256, 0, st>>>(mask, n, css_make_drop(p, seed, site));
    return css_check_launch("css_dropout_mask");
}
