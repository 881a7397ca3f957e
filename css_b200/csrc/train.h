// Internal launch interface of the training-step kernels (train_ops.cu), used by conformer_train.cu.
#pragma once
#include "dropout.h"
#include "gemm.h"
#include <cuda_runtime.h>

// x fp32 (rows, cols; pitch ld) ->
//   out  : bf16 split planes [P][rows][out->ld], columns [cols, out->ld) zero filled              (optional)
//   outT : bf16 split planes of the transpose [P][cols][rows_p], columns [rows, rows_p) zero filled (optional)
//   colsum[c] += cs_scale * sum_r x[r][c]                                                          (optional)
int split_t_launch(const float* x, long rows, int cols, int ld, const PlaneMat* out, const PlaneMat* outT, long rows_p,
                   float* colsum, float cs_scale, cudaStream_t st);
// bf16 planes [P][rows][in.ld] (first `cols` columns) -> [P][cols][rows_p], columns [rows, rows_p) zero filled
int transpose_planes_launch(const PlaneMat& in, long rows, int cols, const PlaneMat& outT, long rows_p, cudaStream_t st);

struct LnBwdArgs {
    const float* x;          // LayerNorm input (M, 256)
    const float* gamma;
    const float* beta;       // GLU variant only
    const float* dn;         // gradient w.r.t. the LayerNorm output (GLU variant: w.r.t. the gated output u)
    const float* relu_out;   // optional forward output of a fused ReLU: dn is masked where it is <= 0
    float* dx;               // (M, 256); may alias dn
    int accumulate;          // 1: dx += ..., 0: dx = ...
    float* dgamma;
    float* dbeta;
    // GLU variant (conv module: u = (n w0 + b0) * sigmoid(n w1 + b1), css/models/conformer.py:293-296)
    const float* pw1_w;      // (2)
    const float* pw1_b;      // (2)
    float* dpw1_w;
    float* dpw1_b;
};
int ln_bwd_launch(const LnBwdArgs& a, long M, int d, float eps, bool glu, cudaStream_t st);

// relative-position attention of the training step on the warp-level tensor cores (attention_train.cu), T <= 256:
// qkv (N*T, 3*H*64) -> out (N*T, H*64); `probs` keeps the attention probabilities for the backward.
// split3: three-term TF32 split (fp32-class accuracy) instead of single TF32 products.
size_t attention_train_scratch_floats(long N, int T, int H);
size_t attention_train_probs_floats(long N, int T, int H);
// drop != null: dropout on the attention probabilities (conformer.py:249); probs keeps the un-dropped probabilities
int attention_train_fwd_launch(const float* qkv, const float* pe, int pe_rows, float* out, float* probs, float* scratch, long N,
                               int T, int H, int dk, bool split3, const CssDrop* drop, cudaStream_t st);
// dO (N*T, H*64) -> dqkv (N*T, 3*H*64) overwritten, dpe (pe_rows, 64) accumulated
int attention_train_bwd_launch(const float* qkv, const float* pe, int pe_rows, const float* probs, const float* dO, float* dqkv,
                               float* dpe, float* scratch, long N, int T, int H, int dk, bool split3, const CssDrop* drop,
                               cudaStream_t st);

struct ConvTrainParams {
    const float *ln_g, *ln_b, *pw1_w, *pw1_b, *dw_w, *dw_b, *bn_g, *bn_b, *pw2_w, *pw2_b;
    float *running_mean, *running_var;     // updated with momentum 0.1 (torch.nn.BatchNorm1d defaults)
};
struct ConvTrainGrads {
    float *ln_g, *ln_b, *pw1_w, *pw1_b, *dw_w, *dw_b, *bn_g, *bn_b, *pw2_w, *pw2_b;
};
// out = x + ConvModule_train(x): batch-statistics BatchNorm.  cpre (M,256) pre-BN activations and bnsave (512: mean, rstd)
// are kept for the backward; stats (512 doubles) is scratch.
// drop != null: out = x + dropout(ConvModule_train(x)) (conformer.py:302), element index = row * 256 + channel
int conv_train_fwd_launch(const float* x, float* out, float* cpre, float* bnsave, double* stats, long N, int T,
                          const ConvTrainParams& p, const CssDrop* drop, cudaStream_t st);
// g (M,256): gradient w.r.t. the module output on entry, w.r.t. its input on return.  du: scratch (M,256); bstats: 512 floats.
// gbranch: gradient w.r.t. the module's BRANCH output (== g without dropout, g * mask / (1-p) with it); read only.
int conv_train_bwd_launch(const float* x, const float* cpre, const float* bnsave, float* g, const float* gbranch, float* du,
                          float* bstats, long N, int T, const ConvTrainParams& p, const ConvTrainGrads& gr, cudaStream_t st);

// y_pred[n][k][t][f] = head[(n,t)][k*nb + f] * feat[(n,t)][f]   (css/models/conformer.py:103-104 with f_orig[..., :nb])
int ypred_launch(const float* head, int ldh, const float* feat, int ldf, float* ypred, long N, int T, int nb, int K,
                 cudaStream_t st);
// dhead[(n,t)][k*nb+f] = head > 0 ? gy[n][k][t][f] * feat[(n,t)][f] : 0 for k < K; the remaining columns of the pitch are zero
int head_bwd_launch(const float* head, int ldh, const float* feat, int ldf, const float* gy, float* dhead, long N, int T,
                    int nb, int K, cudaStream_t st);
