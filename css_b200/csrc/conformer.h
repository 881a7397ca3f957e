// Internal launch interface of the Conformer / BLSTM building-block kernels.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

int layernorm_launch(const float* x, const float* g, const float* b, float* y, long M, int d, float eps, int relu,
                     cudaStream_t st);
int layernorm_planes_launch(const float* x, const float* g, const float* b, float* y, long M, int d, float eps, int relu,
                            void* planes, long plane_stride, int n_planes, cudaStream_t st);
// y = LN(x; g, b); planes = split planes of LN(y; g2, b2) (g2 == null: of y)
int layernorm2_planes_launch(const float* x, const float* g, const float* b, float* y, const float* g2, const float* b2, long M,
                             int d, float eps, int relu, void* planes, long plane_stride, int n_planes, cudaStream_t st);
int attention_launch(const float* qkv, const float* pe, int pe_rows, float* out, long N, int T, int H, int dk,
                     cudaStream_t st);

// warp-level tensor-core attention (attention_tc.cu): T <= 160, dk = 64; fp32 and/or bf16 plane output
bool attention_tc_supported(int T, int dk);
int attention_tc_launch(const void* qkv_planes, long qkv_plane_stride, int in_planes, const void* pe_planes, long pe_plane_stride,
                        int pe_rows, float* out, void* planes, long plane_stride, int n_planes, long N, int T, int H, int dk,
                        cudaStream_t st);

// tcgen05 / TMEM attention (attention_umma.cu): T == 150, dk == 64, two input planes, relative positions
bool attention_umma_supported(int T, int dk, int in_planes, bool has_pe, int pe_rows);
int attention_umma_launch(const void* qkv_planes, long qkv_plane_stride, const void* pe_planes, long pe_plane_stride, int pe_rows,
                          float* out, void* planes, long plane_stride, int n_planes, long N, int H, cudaStream_t st);

struct ConvModuleParams {
    const float* ln_g;      // (d)
    const float* ln_b;      // (d)
    float pw1_w0, pw1_b0;   // pw_conv_1 channel 0 (value path)     -- scalar affine (Conv2d(1,2,1))
    float pw1_w1, pw1_b1;   // pw_conv_1 channel 1 (gate path)
    const float* dw_w;      // (d, ksize) depthwise taps
    const float* dw_b;      // (d)
    const float* bn_scale;  // (d) gamma / sqrt(running_var + eps)
    const float* bn_shift;  // (d) beta - running_mean * scale
    float pw2_w, pw2_b;     // pw_conv_2 scalar affine
    int ksize;
    const float* dw_wt = nullptr;   // optional: the taps transposed to (ksize = 33, d) by conv_taps_transpose_launch -> the
                                    // two-CTAs-per-SM shape of the kernel
};
int conv_taps_transpose_launch(const float* w /* (256, 33) */, float* wt /* (33, 256) */, cudaStream_t st);
// out = x + ConvModule(x).  x, out: (N, T, 256), out must not alias x.  ksize odd, <= 33.
int conv_module_launch(const float* x, float* out, long N, int T, int d, const ConvModuleParams& p, cudaStream_t st);

// (N*T, ldm) rows [c*nb + f] -> masks (N, nb, T, C)
int masks_layout_launch(const float* m, int ldm, float* out, long N, int T, int nb, int C, cudaStream_t st);

// (x - mean_T) / (std_T + eps) over time, emitted as bf16 split planes (features.cu)
int mvn_time_planes_launch(const float* x, long n, int T, int F, int ld, void* planes, long plane_stride, int ldp, int n_planes,
                           cudaStream_t st);

// spectrum -> features -> input MVN -> bf16 planes + magnitude in one kernel (features.cu); feat may be NULL
bool features_planes_supported(int T, int n_pairs);
int features_planes_launch(const void* spec, long stride_d, long stride_chunk, long n_chunks, int T, const int* ipd_l,
                           const int* ipd_r, int n_pairs, int shared_right, float* mag, void* planes, long plane_stride, int ldp, int n_planes,
                           float* feat, int ldf, cudaStream_t st);
