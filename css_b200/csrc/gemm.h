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
    // training (tcgen05 path only)
    const void* relu_mask = nullptr;   // bf16 plane 0 of the forward activation: result zeroed where it is <= 0
    int ldmask = 0;
    int accumulate = 0;                // 1: C += result (split-K over the SMs, fp32 atomics); bias/act/residual/planes unsupported
    // fused LayerNorm (tcgen05 CTA-pair path, N == 256): C gets the GEMM result, `out` the split planes of
    // LayerNorm(C row; ln_g, ln_b, eps 1e-5) -- see gemm_tc_ln_fusable()
    const float* ln_g = nullptr;
    const float* ln_b = nullptr;
    // ln_to_c != 0: C gets LayerNorm(result row; ln_g, ln_b) instead of the raw result (a block's output norm, conformer.py:356),
    // and `out` the split planes of that C -- or, with ln2_g / ln2_b set, of a SECOND LayerNorm of it (the next block's FFN-in norm)
    int ln_to_c = 0;
    const float* ln2_g = nullptr;
    const float* ln2_b = nullptr;
    // MN-major operands (tcgen05 path, training kernels): the operand is stored with its NON-contracted index contiguous,
    //   a_mn: A given as planes of A^T, i.e. a (K, M) row-major matrix (pitch A.ld >= M)  -- dW = dY^T X reads dY as it is
    //   b_mn: W given as planes of W^T, i.e. a (K, N) row-major matrix (pitch W.ld >= N)  -- dX = dY W reads W as it is
    // so neither gradient GEMM needs a transposed copy of an activation or of a weight.
    int a_mn = 0;
    int b_mn = 0;
    long k_valid = 0;   // rows of the MN-major operands that exist in memory (0: K); rows in [k_valid, K) read as zero (TMA fill)
};

int gemm_simt_launch(const float* A, int lda, const float* W, int ldw, float* C, int ldc, long M, int N, int K,
                     GemmEpilogue ep, cudaStream_t st);

// ---- tcgen05/TMEM path (gemm_tc.cu) ----
// A matrix stored as `planes` bf16 split-float planes: element (p, r, k) at ptr[p*plane_stride + r*ld + k].
struct PlaneMat {
    void* ptr;
    long plane_stride;   // elements
    int ld;              // elements, multiple of 8
    int planes;
};
// C (fp32, optional) and/or `out` planes (optional) = epilogue(A . W^T) with P-plane operands
// (P = 1 bf16, 2 bf16x3, 3 bf16x6).  K must be a multiple of 64.
int gemm_tc_planes(const PlaneMat& A, const PlaneMat& W, int P, long M, int N, int K, const GemmEpilogue& e, float* C, int ldc,
                   const PlaneMat* out, cudaStream_t st);
int split_planes_launch(const float* x, long rows, int cols, int ld, const PlaneMat& out, cudaStream_t st);
int sm_count();
// true when gemm_tc_planes can apply GemmEpilogue::ln_g / ln_b in its epilogue for this shape (otherwise run LayerNorm separately)
bool gemm_tc_ln_fusable(long M, int N, int P);
