/* css_b200 -- C ABI of the B200-native CSS hot path (libcss_b200.so).
 *
 * The reference (desh2608/css) is pure Python: it has no FFI boundary of its own.  Its plug-in
 * points for this path are Python objects; each entry point below is what the
 * Python class/function named in its comment binds through ctypes (css_b200/_native.py).
 *
 * Conventions: every pointer is a DEVICE pointer unless the name ends in _host; the caller owns
 * and allocates all buffers; calls are stream-ordered on `stream` and never synchronise; the
 * return value is 0 on success, negative on error with the text in css_last_error().
 * "KP" = 264 is the padded bin pitch of bin-fastest arrays (257 bins of the 512-point STFT).
 */
#ifndef CSS_B200_H
#define CSS_B200_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* css_stream_t;

const char* css_last_error(void);
int css_version(void);
unsigned long long css_launch_count(int reset);   /* kernels launched since last reset */
/* optional CUDA-event timing per kernel category: 0 gemm, 1 attention, 2 conv module, 3 layernorm, 4 lstm step */
void css_profile_enable(int on);
int css_profile_collect(double* ms_host, double* work_host, long* count_host, int n_cat);

/* ---- STFT / features: css/executor/feature.py:45-82 (STFT.forward), :105-134, :156-175 ---- */
/* spec[sig][frame][KP] complex64; frame g covers samples [256 g, 256 g + 512), Hann(512). */
int css_stft(const float* x, long n_sig, long sig_stride, long n_frames, void* spec, css_stream_t stream);
/* same, into a frame range of a larger buffer: signal s is read at x + s*sig_stride and written at spec + s*spec_stride frames */
int css_stft_range(const float* x, long n_sig, long sig_stride, long n_frames, void* spec, long spec_stride, css_stream_t stream);
/* centred-STFT frame 0 (reflect pad) of every chunk: css/executor/beamformer.py:137-145 */
int css_stft_edge(const float* x, long n_sig, long sig_stride, long n_chunks, long chunk_hop, void* edge,
                  css_stream_t stream);
/* mag (B,257,T), feat (B,T,ldf) = [MVN(mag ch0) | IPD pairs]; spec element (d,b,t,k) at
 * spec[(d*stride_d + b*stride_chunk + t)*KP + k] */
int css_features(const void* spec, long stride_d, long stride_chunk, long n_chunks, int T, const int* ipd_l,
                 const int* ipd_r, int n_pairs, float* mag, float* feat, int ldf, css_stream_t stream);
/* css/models/conformer.py:95, css/models/blstm.py:80 */
int css_mvn_time(float* x, long n, int T, int F, int ld, css_stream_t stream);

/* ---- dense layers: torch.nn.Linear call sites css/models/conformer.py:79-81,124,205-209,314-320 ---- */
/* C = residual + alpha * act(A . W^T + bias); act 0 none / 1 ReLU; clamp_max > 0 caps after act */
int css_gemm_f32(const float* A, int lda, const float* W, int ldw, const float* bias, float* C, int ldc, long M,
                 int N, int K, int act, float alpha, const float* residual, int ldr, float clamp_max,
                 css_stream_t stream);
int css_layernorm(const float* x, const float* g, const float* b, float* y, long M, int d, float eps, int relu,
                  css_stream_t stream);
/* css/models/conformer.py:229-250 (MultiHeadedAttention core with relative position keys) */
int css_relpos_attention(const float* qkv, const float* pe, int pe_rows, float* out, long N, int T, int H, int dk,
                         css_stream_t stream);
/* same contract on the warp-level tensor-core path (mma.sync bf16, three-term split, fp32 accumulate); T <= 160 */
int css_relpos_attention_tc(const float* qkv, const float* pe, int pe_rows, float* out, long N, int T, int H, int dk,
                            css_stream_t stream);
/* css/executor/separator.py:100-104: (N*T, C*nb) -> (N, nb, T, C) */
int css_masks_layout(const float* m, int ldm, float* out, long N, int T, int nb, int C, css_stream_t stream);

/* ---- Conformer mask estimator: css/models/conformer.py:83-104 (Conformer.forward) ---- */
typedef struct {
    const float *ln_g, *ln_b, *w1, *b1, *w2, *b2;        /* FeedForward: LN, Linear(d,dffn), Linear(dffn,d) */
} css_ffn_weights;
typedef struct {
    css_ffn_weights ff_in;
    const float *att_ln_g, *att_ln_b, *wqkv, *bqkv, *wo, *bo;   /* wqkv = cat(linear_q,k,v) (3d,d) */
    const float *conv_ln_g, *conv_ln_b, *dw_w, *dw_b, *bn_scale, *bn_shift;   /* dw_w centre-padded to 33 taps */
    float pw1_w0, pw1_b0, pw1_w1, pw1_b1, pw2_w, pw2_b;         /* scalar "pointwise" affines (host values) */
    css_ffn_weights ff_out;
    const float *out_ln_g, *out_ln_b;
} css_conformer_block;
typedef struct {
    int idim, ldin, d_model, heads, d_ffn, num_blocks, num_bins, num_out, pe_rows;
    const float *embed_w, *embed_b, *embed_ln_g, *embed_ln_b;   /* embed_w (d, ldin), zero padded */
    const float* pe;                                            /* (pe_rows, d/heads) or NULL */
    const float *head_w, *head_b;                               /* (num_bins*num_out, d) */
} css_conformer_desc;

void* css_conformer_create(const css_conformer_desc* desc_host, const css_conformer_block* blocks_host);
void css_conformer_destroy(void* model);
size_t css_conformer_workspace_bytes(const void* model, long N, int T);
/* feat (N*T, ldin) is normalised IN PLACE when apply_mvn != 0.  Exactly one of masks (N,nb,T,C;
 * clamped to <= clamp_max when clamp_max > 0) or head (N*T, ld_head; ReLU only) may be non-NULL. */
int css_conformer_forward(void* model, float* feat, long N, int T, int apply_mvn, float clamp_max, float* masks,
                          float* head, int ld_head, void* workspace, size_t workspace_bytes, css_stream_t stream);

/* FeatureExtractor.forward (css/executor/feature.py:156-175) + the model's input MVN (conformer.py:95) + Conformer.forward
 * (conformer.py:83-104) in one call, as css/executor/separator.py:86-104 chains them: the fp32 feature matrix is never
 * materialised.  spec/strides/ipd tables as css_features; mag (N,257,T) is css_features' magnitude output.  Tensor-core
 * GEMM modes only; css_features_planes_supported(T, n_pairs) says whether the fused front end takes the shape.
 * shared_right != 0 is the caller's claim that every pair has the same right channel (ipd_r[p] == ipd_r[0], as in
 * conf/config_7ch.yaml "1,0;2,0;..."): that channel is then loaded once per frame; the kernel checks the claim and traps
 * if it is false.  0 is always correct. */
int css_conformer_forward_spec(void* model, const void* spec, long stride_d, long stride_chunk, const int* ipd_l,
                               const int* ipd_r, int n_pairs, int shared_right, long N, int T, float clamp_max, float* mag, float* masks,
                               float* head, int ld_head, void* workspace, size_t workspace_bytes, css_stream_t stream);
int css_features_planes_supported(int T, int n_pairs);
/* The front end of that call on its own: magnitude (N,257,T) + the n_planes bf16 split planes (n_planes x (N*T, ldp),
 * plane_stride elements apart, K padding zeroed) of the time-normalised features; feat (N*T, ldf) fp32 = css_features'
 * feature output, or NULL. */
int css_features_planes(const void* spec, long stride_d, long stride_chunk, long n_chunks, int T, const int* ipd_l,
                        const int* ipd_r, int n_pairs, int shared_right, float* mag, void* planes, long plane_stride, int ldp, int n_planes,
                        float* feat, int ldf, css_stream_t stream);

/* ---- BLSTM mask estimator: css/models/blstm.py:68-88 (BLSTM.forward) ---- */
typedef struct {
    const float *w_ih;      /* (4*hdim, hdim) = cat(weight_ih_l0, weight_ih_l0_reverse) */
    const float *b_gates;   /* (4*hdim) = cat(bias_ih+bias_hh fwd, bias_ih+bias_hh reverse) */
    const float *w_hh_f, *w_hh_b;   /* (2*hdim, hdim/2) each */
    const float *ln_g, *ln_b;
} css_blstm_layer;
typedef struct {
    int idim, ldin, hdim, num_layers, num_bins, num_out;
    const float *embed_w, *embed_b, *embed_ln_g, *embed_ln_b;
    const float *head_w, *head_b;
} css_blstm_desc;
void* css_blstm_create(const css_blstm_desc* desc_host, const css_blstm_layer* layers_host);
void css_blstm_destroy(void* model);
size_t css_blstm_workspace_bytes(const void* model, long N, int T);
int css_blstm_forward(void* model, float* feat, long N, int T, int apply_mvn, float clamp_max, float* masks,
                      float* head, int ld_head, void* workspace, size_t workspace_bytes, css_stream_t stream);

/* ---- mask post-processing + stitching: css/executor/separator.py:165-250, css/executor/stitcher.py ---- */
/* svT (30,7,hi-lo) complex64; like (B,2,30); idx (B,3) = {argmax spk0, argmax spk1, merged?}.  masks (B,257,T,C)
 * are updated in place (output 0 overwritten with 1e-12 where the two DOAs are within thresh degrees). */
int css_doa_merge(const void* spec, long stride_d, long stride_chunk, long n_chunks, int T, const void* svT,
                  const float* angles, int lo, int hi, float thresh, float* masks, int C, float* like, int* idx,
                  css_stream_t stream);
/* Stitcher.get_stitch (stitcher.py:14-52): swap[n]=1 <=> perm [1,0]; sims (B-1,4) optional */
int css_stitch_perm(const float* mags, const float* masks, long n_chunks, int T, int C, int margin, int* swap,
                    float* sims, css_stream_t stream);
/* stitcher.py:55-60: state[0]=0, state[n+1]=state[n]^swap[n] */
int css_perm_states(const int* swap, long n_chunks, int* state, css_stream_t stream);
/* Stitcher.get_connect (stitcher.py:62-110): out [3][L][KP], L = hop*(B-1)+T */
int css_stitch_connect(const float* masks, const int* state, long n_chunks, int T, int C, int hop, float* out,
                       css_stream_t stream);

/* ---- MVDR beamformer: css/executor/beamformer.py:126-183 + asteroid compute_scm / SoudenMVDRBeamformer ---- */
/* spec [7][frames][KP], edge [7][B][KP], stitched [3][L][KP] -> Y [B][2][T][KP] complex64,
 * energy [B][2][css_mvdr_energy_blocks()][2] partial sums (|y|^2, |mask x0|^2) */
int css_mvdr(const void* spec, const void* edge, long frames, long n_chunks, const float* stitched, long L,
             int hop_frames, int T, void* Y, float* energy, css_stream_t stream);
int css_mvdr_energy_blocks(void);
/* normalize_scale (beamformer.py:165-183) + dedup (beamformer.py:68-80), in place; scal (B,5) */
int css_mvdr_finalize(void* Y, const float* energy, long n_chunks, int T, int rescale, int dedup, float* scal,
                      css_stream_t stream);
/* Beamformer.normalize_scale as a stand-alone call (beamformer.py:165-183): result, ref (n_chunks, n) complex64, mask
 * (n_chunks, n) fp32 -> out = result * rms(mask * ref) / rms(result) per chunk */
int css_normalize_scale(const void* result, const void* ref, const float* mask, long n_chunks, long n, void* out,
                        css_stream_t stream);
/* torch.istft + placement (beamformer.py:83-118): plan (n_items,4) int32 {chunk, s_begin, s_end, dst};
 * out (2, out_len) is accumulated into; items of one call must not overlap in out */
int css_istft_place(const void* Y, int T, const int* plan, int n_items, float* out, long out_len, css_stream_t stream);
/* beamformer.py:122-123; scratch: `rows` uint32 */
int css_peak_normalize(float* x, int rows, long n, float target, unsigned* scratch, css_stream_t stream);
/* No counterpart in the reference.  css_peak_normalize for callers that
 * cannot allocate (CUDA graph capture, serving loops): the per-row peak bits live in a 256-byte file-scope device array
 * inside the library, so there is no scratch argument and the call performs no allocation of any kind.
 * rows <= css_peak_normalize_noalloc_max_rows() (64), otherwise -1 with a message.  Every call shares that array
 * (cleared at the start of each call, stream-ordered): calls that overlap on one GPU race. */
int css_peak_normalize_noalloc(float* x, int rows, long n, float target, css_stream_t stream);
int css_peak_normalize_noalloc_max_rows(void);

/* The two halves of css_peak_normalize for a recording spread over several GPUs: bits[r] = max(bits[r], bits of
 * max|x[r][:n]|) (caller zeroes bits; combine across ranks with all-reduce MAX), then x[r][:n] = x * target / peak.
 * x is (rows, ld), n <= ld. */
int css_absmax(const float* x, int rows, long n, long ld, unsigned* bits, css_stream_t stream);
int css_peak_scale(float* x, int rows, long n, long ld, float target, const unsigned* bits, css_stream_t stream);

/* ---- PIT-MSE: css/objectives/pit.py:5-62, css/objectives/mse.py:29-37 ---- */
/* est/tgt (B,K,n); pair_ws (B,K,K) f64; loss_b (B); perm (B,K) int32; loss (1); grad (B,K,n) or NULL */
int css_pit_mse(const float* est, const float* tgt, long B, int K, long n, double* pair_ws, float* loss_b, int* perm,
                float* loss, float* grad, float gscale, css_stream_t stream);

/* ---- training step: css/trainer/iteration_types.py:34-63 (train_one_epoch body), css/objectives/mse.py:22-37 ---- */
/* All parameters live in ONE flat fp32 buffer and all gradients in a second buffer of the same layout (the torch
 * Parameters of css_b200.models.Conformer are views into them); the structs below give element offsets.  Within a
 * block, linear_q/k/v weights (and biases) must be adjacent in that order: wqkv/bqkv name the first of the three. */
typedef struct { long ln_g, ln_b, w1, b1, w2, b2; } css_ffn_offsets;
typedef struct {
    css_ffn_offsets ff_in;
    long att_ln_g, att_ln_b, wqkv, bqkv, wo, bo;
    long conv_ln_g, conv_ln_b, pw1_w, pw1_b, dw_w, dw_b, bn_g, bn_b, pw2_w, pw2_b;
    css_ffn_offsets ff_out;
    long out_ln_g, out_ln_b;
} css_conformer_block_offsets;
typedef struct {
    int idim, d_model, heads, d_ffn, num_blocks, num_bins, num_out, num_spk, pe_rows, ksize;
    long embed_w, embed_b, embed_ln_g, embed_ln_b, pe /* -1: no relative positions */, head_w, head_b;
} css_conformer_train_desc;
/* params/grads: flat buffers; bn_running: [num_blocks][2][d_model] running mean / var (updated in place, momentum 0.1) */
void* css_trainer_create(const css_conformer_train_desc* desc_host, const css_conformer_block_offsets* blocks_host,
                         float* params, float* grads, float* bn_running);
void css_trainer_destroy(void* trainer);
/* nn.Dropout(p) at the six sites of css/models/conformer.py:127,249,254,302,317,319 for the following forward/backward
 * pairs; masks are a pure function of (seed, site, element) -- pass a new seed per step.  p = 0 (default) disables it. */
int css_trainer_set_dropout(void* trainer, float p, unsigned seed);
/* the keep-mask (uint8, 1 = kept) the training kernels apply at (p, seed, site) to elements 0..n-1: site 0 = embedding,
 * block l: 1 + 8 l + k with k = 0 FFN-in hidden, 1 FFN-in output, 2 attention probabilities ([head][n][i][T rounded up to
 * 32]), 3 attention output, 4 conv module, 5 FFN-out hidden, 6 FFN-out output; 2-D sites index row * cols + col */
int css_dropout_mask(unsigned char* mask, long n, float p, unsigned seed, unsigned site, css_stream_t stream);
size_t css_trainer_workspace_bytes(const void* trainer, long N, int T);
/* Conformer.forward in train mode (css/models/conformer.py:83-104; dropout probability 0): feat (N,T,idim) raw features
 * -> y_pred (N,num_spk,T,num_bins) and optionally the head activations masks_head (N*T, ld_head).  Activations needed by
 * the backward stay in `workspace`, which must be passed unchanged to css_trainer_backward. */
int css_trainer_forward(void* trainer, const float* feat, long N, int T, float* y_pred, float* masks_head, int ld_head,
                        void* workspace, size_t workspace_bytes, css_stream_t stream);
/* d loss / d y_pred (N,num_spk,T,num_bins) -> gradients ACCUMULATED into the flat gradient buffer. */
int css_trainer_backward(void* trainer, const float* feat, const float* grad_y_pred, long N, int T, void* workspace,
                         size_t workspace_bytes, css_stream_t stream);
/* The same backward in pieces, so that the caller can start the gradient all-reduce of finished layers while earlier
 * layers are still running: parts are num_blocks (mask head), l (encoder block l) and -1 (input embedding), processed
 * from part_hi down to part_lo.  The whole range num_blocks..-1 must be covered in descending order. */
int css_trainer_backward_part(void* trainer, const float* feat, const float* grad_y_pred, long N, int T, int part_hi,
                              int part_lo, void* workspace, size_t workspace_bytes, css_stream_t stream);
/* torch.nn.utils.clip_grad_norm_ + torch.optim.Adam(weight_decay) over flat buffers (scripts/python/train.py:209-211) */
int css_grad_sqnorm(const float* g, long n, double* out, css_stream_t stream);
int css_adam_step(float* p, const float* g, float* m, float* v, long n, float lr, float beta1, float beta2, float eps,
                  float weight_decay, long step, float gscale, const double* sqnorm, float max_norm, float* norm_out,
                  css_stream_t stream);

/* ---- on-the-fly training mixer: css/datasets/separation.py:163-267 ---- */
/* One output window: utterance 1 (pool offset/length per channel) starts at mixture sample 0, utterance 2 at start2; the
 * window covers mixture samples [win_start, win_start + W), zero outside the utterances (lhotse pad + mix + window). */
typedef struct { long off1, len1, off2, len2, start2, win_start; } css_mix_item;
/* pool: C channel planes, ch_stride samples apart.  mix (B,C,W); s1, s2 (B,W) = channel 0 of the padded utterances */
int css_mix_windows(const float* pool, long ch_stride, int C, const void* items, int B, int W, float* mix, float* s1,
                    float* s2, css_stream_t stream);
/* x (B,C,W) += g * noise[(off + n) mod noise_len] with g from the channel-0 powers and snr_db[b]; snr_db[b] = NaN: skip */
int css_mix_add_noise(float* x, int B, int C, int W, const float* noise, long noise_ch_stride, long noise_len,
                      const long* noise_off, const float* snr_db, css_stream_t stream);
/* y[b] = energy-normalised FIR(x[b], rirs[rir_idx[b]]) for rir_idx[b] >= 0 (other windows are left unwritten) */
int css_mix_reverb(const float* x, float* y, int B, int C, int W, const float* rirs, int taps, const int* rir_idx,
                   double* energy, css_stream_t stream);

/* ---- Conv-TasNet: css/models/conv_tasnet.py:17-330.  Activations time-major (n*T, C); every
 * Conv1d(kernel 1) is css_gemm_f32; these cover the rest ---- */
/* encoder Conv1d(1, N, L, stride): x (n, len) -> out (n*T, N), T = (len - L) / stride + 1 */
int css_ctn_encoder(const float* x, long n, long len, const float* w, const float* b, int N, int L, int stride, float* out,
                    css_stream_t stream);
/* nn.LayerNorm over the C channels of every row ("cln") */
int css_row_layernorm(const float* x, const float* g, const float* b, long M, int C, float eps, float* y, css_stream_t stream);
/* global layer norm of prelu(x) ("gln"): stats (n, 2) f64 = {sum, sum of squares} per sample; apply; and the dilated
 * depthwise conv that reads gLN(prelu(x)) on the fly (zero padding on the normalised signal, odd kernel, non-causal) */
int css_ctn_prelu_stats(const float* x, long n, long per_sample, const float* alpha, double* stats, css_stream_t stream);
int css_ctn_norm_apply(const float* x, long n, long T, int C, const float* alpha, const double* stats, const float* g,
                       const float* b, float eps, float* y, css_stream_t stream);
int css_ctn_dwconv(const float* x, long n, long T, int C, int P, int dil, const float* alpha, const double* stats,
                   const float* g, const float* b, float eps, const float* w, const float* wb, float* y, css_stream_t stream);
/* masks x encoder output -> ConvTranspose1d(N, 1, L, stride) -> y (S, n, (T-1) stride + L); frames: S n T L floats scratch */
int css_ctn_decode(const float* enc, const float* masks, int ldm, long n, long T, int N, int L, int stride, int S,
                   const float* wd, const float* bias, float* frames, float* y, css_stream_t stream);

/* tcgen05/TMEM GEMM from fp32 operands (same contract as css_gemm_f32): operands are split into `planes` bf16 planes
 * in `scratch` (1 = bf16, 2 = bf16x3, 3 = bf16x6 product terms), accumulation in fp32 */
size_t css_gemm_tc_scratch_bytes(long M, int N, int K, int planes);
int css_gemm_tc(const float* A, int lda, const float* W, int ldw, const float* bias, float* C, int ldc, long M, int N,
                int K, int act, float alpha, const float* residual, int ldr, float clamp_max, int planes, void* scratch,
                size_t scratch_bytes, css_stream_t stream);

/* MN-major operand forms of the tcgen05 GEMM (the gradient GEMMs of the training step read dY, X and W as they are stored,
 * no transposed copies): mode 1: C = A . W^T with W given as Wt = W^T (K, N); mode 3: A given as At = A^T (K, M) as well and
 * the result ACCUMULATED into C (split-K, fp32 atomics).  Test / tooling entry; K % 64 == 0, N > 128. */
int css_gemm_tc_mn(const float* A, int lda, const float* Wt, int ldw, float* C, int ldc, long M, int N, int K, int mode,
                   int planes, void* scratch, size_t scratch_bytes, css_stream_t stream);

/* GEMM engine of the mask networks: 0 = fp32 SIMT, 1 = tcgen05 bf16x3 split (fp32-class accuracy), 2 = tcgen05 bf16 */
int css_set_gemm_mode(int mode);
int css_get_gemm_mode(void);

/* Launch plan of the tensor-core BLSTM recurrence (blstm.py:156-160): 0 = chosen from (sequences, hidden size); 1 = one launch
 * per time step with operands resident in shared memory; 3 = all T steps of a layer in one cooperative launch (<= 80
 * sequences, falls back to 0 when the grid cannot be co-resident); 12 / 22 = streamed 32- / 64-sequence CTAs per step.
 * Initialised from CSS_B200_LSTM_TILE.  Plans 1 and 3 compute identical bits; this exists for A/B timing and the parity test. */
int css_set_lstm_plan(int plan);
int css_get_lstm_plan(void);

#ifdef __cplusplus
}
#endif
#endif
