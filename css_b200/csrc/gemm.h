// Internal GEMM interface shared by the SIMT fp32 path and the tcgen05 path.
#pragma once
#include <cuda_runtime.h>

struct GemmEpilogue {
    const float* bias;       // per output column, may be null
    int act;                 // 0 none, 1 ReLU
    float alpha;             // scale applied after the activation
    const float* residual;   // added last, may be null (may alias C)
    int ldr;
    float clamp_max;         // > 0: min(v, clamp_max) after the activation
};

int gemm_simt_launch(const float* A, int lda, const float* W, int ldw, float* C, int ldc, long M, int N, int K,
                     GemmEpilogue ep, cudaStream_t st);

// tcgen05/TMEM path (gemm_tc.cu).  mode 1 = bf16x3 split (fp32-accurate), 2 = plain bf16.
// Returns GEMM_TC_UNSUPPORTED when the shape cannot be handled (caller falls back to SIMT fp32).
#define GEMM_TC_UNSUPPORTED 1000
int gemm_tc_launch(const float* A, int lda, const float* W, int ldw, float* C, int ldc, long M, int N, int K,
                   GemmEpilogue ep, int mode, cudaStream_t st);
