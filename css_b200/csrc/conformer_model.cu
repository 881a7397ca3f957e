// Conformer mask estimator forward (css/models/conformer.py:83-104,152-165,348-362), orchestrated
// natively: one C call launches the whole 16-block network on the caller's stream.
#include "common.cuh"
#include "conformer.h"
#include "gemm.h"
#include "../../include/css_b200.h"
#include <cstdlib>
#include <new>
#include <vector>

static int g_gemm_mode = 1;   // tcgen05 bf16x3: fp32-class accuracy (masks within 1e-4 of the fp32 oracle)
extern "C" int css_set_gemm_mode(int mode) {
    CSS_REQUIRE(mode >= 0 && mode <= 2, "css_set_gemm_mode: unknown mode %d", mode);
    g_gemm_mode = mode;
    return 0;
}
extern "C" int css_get_gemm_mode(void) { return g_gemm_mode; }

int gemm_dispatch(const float* A, int lda, const float* W, int ldw, float* C, int ldc, long M, int N, int K,
                  GemmEpilogue ep, cudaStream_t st) {
    css_prof_begin(0, 2.0 * (double)M * N * K, st);
    int rc = gemm_simt_launch(A, lda, W, ldw, C, ldc, M, N, K, ep, st);
    css_prof_end(st);
    return rc;
}

// bf16 split planes of every Linear weight (tcgen05 path), built lazily on the first tensor-core forward
struct BlockPlanes { PlaneMat ff_in_w1, ff_in_w2, wqkv, wo, ff_out_w1, ff_out_w2; };
struct ConformerModel {
    css_conformer_desc d;
    std::vector<css_conformer_block> blk;
    void* plane_mem = nullptr;
    PlaneMat embed_w{}, head_w{}, pe{};
    std::vector<BlockPlanes> bp;
    float* taps_t = nullptr;     // [num_blocks][33][256]: depthwise taps transposed for the conv module (inside plane_mem)
};
constexpr int kWeightPlanes = 2;

static int build_weight_planes(ConformerModel* m, cudaStream_t st) {
    if (m->plane_mem) return 0;
    const css_conformer_desc& c = m->d;
    const size_t d = c.d_model, f = c.d_ffn, nh = (size_t)c.num_bins * c.num_out;
    auto bytes = [](size_t rows, size_t k) { return (kWeightPlanes * rows * k * 2 + 255) & ~(size_t)255; };
    const size_t dk = d / c.heads;
    const size_t taps_bytes = ((size_t)c.num_blocks * 33 * d * 4 + 255) & ~(size_t)255;
    size_t total = bytes(d, c.ldin) + bytes(nh, d) + (c.pe ? bytes(c.pe_rows, dk) : 0) + c.num_blocks * (2 * bytes(f, d) + 2 * bytes(d, f) + bytes(3 * d, d) + bytes(d, d)) + taps_bytes;
    cudaError_t e = cudaMalloc(&m->plane_mem, total);
    CSS_REQUIRE(e == cudaSuccess, "conformer: cannot allocate %zu B for the bf16 weight planes: %s", total, cudaGetErrorString(e));
    char* wp = static_cast<char*>(m->plane_mem);
    int rc = 0;
    auto mk = [&](const float* w, size_t rows, size_t k) {
        PlaneMat pm = {wp, (long)(rows * k), (int)k, kWeightPlanes};
        wp += bytes(rows, k);
        if (!rc) rc = split_planes_launch(w, (long)rows, (int)k, (int)k, pm, st);
        return pm;
    };
    m->embed_w = mk(c.embed_w, d, c.ldin);
    m->head_w = mk(c.head_w, nh, d);
    if (c.pe) m->pe = mk(c.pe, c.pe_rows, dk);
    m->bp.resize(c.num_blocks);
    for (int l = 0; l < c.num_blocks; ++l) {
        const css_conformer_block& b = m->blk[l];
        BlockPlanes& q = m->bp[l];
        q.ff_in_w1 = mk(b.ff_in.w1, f, d);
        q.ff_in_w2 = mk(b.ff_in.w2, d, f);
        q.wqkv = mk(b.wqkv, 3 * d, d);
        q.wo = mk(b.wo, d, d);
        q.ff_out_w1 = mk(b.ff_out.w1, f, d);
        q.ff_out_w2 = mk(b.ff_out.w2, d, f);
    }
    m->taps_t = reinterpret_cast<float*>(wp);
    for (int l = 0; l < c.num_blocks && !rc; ++l) rc = conv_taps_transpose_launch(m->blk[l].dw_w, m->taps_t + (size_t)l * 33 * d, st);
    if (rc) { cudaFree(m->plane_mem); m->plane_mem = nullptr; m->taps_t = nullptr; }
    return rc;
}

extern "C" void* css_conformer_create(const css_conformer_desc* desc, const css_conformer_block* blocks) {
    if (!desc || !blocks) { css_set_error("css_conformer_create: null descriptor"); return nullptr; }
    if (desc->d_model != 256 || desc->d_model / desc->heads != 64) {
        css_set_error("css_conformer_create: kernels are built for attention_dim 256 with 64-wide heads (got %d/%d)",
                      desc->d_model, desc->heads);
        return nullptr;
    }
    if (desc->ldin % 16 || desc->ldin < desc->idim || desc->d_ffn % 128 || desc->d_ffn > 1024) {
        css_set_error("css_conformer_create: bad ldin=%d (idim %d) or d_ffn=%d", desc->ldin, desc->idim, desc->d_ffn);
        return nullptr;
    }
    ConformerModel* m = new (std::nothrow) ConformerModel();
    if (!m) { css_set_error("css_conformer_create: out of host memory"); return nullptr; }
    m->d = *desc;
    m->blk.assign(blocks, blocks + desc->num_blocks);
    return m;
}

extern "C" void css_conformer_destroy(void* model) {
    ConformerModel* m = static_cast<ConformerModel*>(model);
    if (m && m->plane_mem) cudaFree(m->plane_mem);
    delete m;
}

static inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }
static inline int head_pitch(const css_conformer_desc& d) { return ((d.num_bins * d.num_out + 7) / 8) * 8; }

extern "C" size_t css_conformer_workspace_bytes(const void* model, long N, int T) {
    const ConformerModel* m = static_cast<const ConformerModel*>(model);
    const size_t M = (size_t)N * T, d = m->d.d_model;
    const size_t wide = (size_t)(m->d.d_ffn > 3 * m->d.d_model ? m->d.d_ffn : 3 * m->d.d_model);
    // fp32 path: x, y, a (M x d), h (M x wide), head;  tensor-core path adds the bf16 plane buffers of the GEMM inputs:
    // LN output, attention output (2 x M x d each), FFN hidden (2 x M x d_ffn), features (2 x M x ldin)
    return 3 * align256(M * d * 4) + align256(M * wide * 4) + align256(M * head_pitch(m->d) * 4) +
           2 * align256(2 * M * d * 2) + align256(2 * M * (size_t)m->d.d_ffn * 2) + align256(2 * M * (size_t)m->d.ldin * 2);
}

static int ffn(const css_ffn_weights& w, float* x, float* y, float* h, long M, int d, int dffn, cudaStream_t st) {
    // x <- x + 0.5 * W2 relu(W1 LN(x) + b1) + b2      (css/models/conformer.py:322-326,350,353)
    int rc = layernorm_launch(x, w.ln_g, w.ln_b, y, M, d, 1e-5f, 0, st);
    if (rc) return rc;
    GemmEpilogue e1 = {w.b1, 1, 1.f, nullptr, 0, 0.f};
    rc = gemm_dispatch(y, d, w.w1, d, h, dffn, M, dffn, d, e1, st);
    if (rc) return rc;
    GemmEpilogue e2 = {w.b2, 0, 0.5f, x, d, 0.f};
    return gemm_dispatch(h, dffn, w.w2, dffn, x, d, M, d, dffn, e2, st);
}

// ---------------------------------------------------------------------------------------------
// Tensor-core forward: every Linear runs on the tcgen05 GEMM (gemm_tc.cu) over bf16 split planes.
// The producers of GEMM inputs (input MVN, LayerNorm, the FFN-in GEMM epilogue, attention) write planes directly.
// ---------------------------------------------------------------------------------------------
static int tc_gemm(const PlaneMat& A, const PlaneMat& W, int P, long M, int N, int K, GemmEpilogue e, float* C, int ldc,
                   const PlaneMat* out, cudaStream_t st) {
    css_prof_begin(0, 2.0 * (double)M * N * K, st);
    int rc = gemm_tc_planes(A, W, P, M, N, K, e, C, ldc, out, st);
    css_prof_end(st);
    return rc;
}

// ln_done: yP already holds the planes of LN(x) (written by the previous block's output norm).
// next_g/next_b: LayerNorm that follows this FFN (the attention-in norm after the first FFN of a block); when the GEMM can
// take it in its epilogue, yP leaves holding the planes of LN(x_new) and *fused is set -- the caller then skips that launch.
// norm_x (the block's second FFN): next_g/next_b is the block's OUTPUT norm -- x itself leaves normalised (conformer.py:356) -- and
// yP holds the planes of that x, or of LN(x; next2_g, next2_b), the following block's FFN-in norm, when next2_g is given.
static int ffn_tc(const css_ffn_weights& w, const PlaneMat& w1, const PlaneMat& w2, float* x, const PlaneMat& yP, const PlaneMat& hP,
                  int P, long M, int d, int dffn, bool ln_done, const float* next_g, const float* next_b, bool* fused, cudaStream_t st,
                  bool norm_x = false, const float* next2_g = nullptr, const float* next2_b = nullptr) {
    int rc = ln_done ? 0 : layernorm_planes_launch(x, w.ln_g, w.ln_b, nullptr, M, d, 1e-5f, 0, yP.ptr, yP.plane_stride, P, st);
    if (rc) return rc;
    GemmEpilogue e1 = {w.b1, 1, 1.f, nullptr, 0, 0.f};
    if ((rc = tc_gemm(yP, w1, P, M, dffn, d, e1, nullptr, 0, &hP, st))) return rc;
    GemmEpilogue e2 = {w.b2, 0, 0.5f, x, d, 0.f};
    const bool fuse = next_g && gemm_tc_ln_fusable(M, d, P);
    if (fused) *fused = fuse;
    if (fuse) {
        e2.ln_g = next_g;
        e2.ln_b = next_b;
        e2.ln_to_c = norm_x ? 1 : 0;
        e2.ln2_g = norm_x ? next2_g : nullptr;
        e2.ln2_b = norm_x ? next2_b : nullptr;
        return tc_gemm(hP, w2, P, M, d, dffn, e2, x, d, &yP, st);
    }
    return tc_gemm(hP, w2, P, M, d, dffn, e2, x, d, nullptr, st);
}

// Input of the fused front end (css_conformer_forward_spec): the chunk spectra instead of a feature matrix.
struct SpecInput {
    const void* spec;
    long stride_d, stride_chunk;
    const int *ipd_l, *ipd_r;
    int n_pairs, shared_right;
    float* mag;
};

static int forward_tc(ConformerModel* m, float* feat, const SpecInput* si, long N, int T, int apply_mvn, float clamp_max, float* masks,
                      float* head, int ld_head, char* wp, int mode, cudaStream_t st) {
    const css_conformer_desc& c = m->d;
    const long M = N * T;
    const int d = c.d_model, dffn = c.d_ffn, nh = c.num_bins * c.num_out, P = (mode == 1) ? 2 : 1;
    const size_t wide = (size_t)(dffn > 3 * d ? dffn : 3 * d);
    float* x = (float*)wp;  wp += align256((size_t)M * d * 4);
    float* y = (float*)wp;  wp += align256((size_t)M * d * 4);
    float* a = (float*)wp;  wp += align256((size_t)M * d * 4);
    float* h = (float*)wp;  wp += align256((size_t)M * wide * 4);
    float* hd = (float*)wp; wp += align256((size_t)M * head_pitch(c) * 4);
    PlaneMat yP = {wp, M * (long)d, d, P};        wp += align256(2 * (size_t)M * d * 2);
    PlaneMat aP = {wp, M * (long)d, d, P};        wp += align256(2 * (size_t)M * d * 2);
    PlaneMat hP = {wp, M * (long)dffn, dffn, P};  wp += align256(2 * (size_t)M * dffn * 2);
    PlaneMat fP = {wp, M * (long)c.ldin, c.ldin, P};
    int rc;
    if ((rc = build_weight_planes(m, st))) return rc;
    if (si)
        rc = features_planes_launch(si->spec, si->stride_d, si->stride_chunk, N, T, si->ipd_l, si->ipd_r, si->n_pairs, si->shared_right, si->mag,
                                    fP.ptr,
                                    fP.plane_stride, c.ldin, P, nullptr, 0, st);
    else if (apply_mvn) rc = mvn_time_planes_launch(feat, N, T, c.idim, c.ldin, fP.ptr, fP.plane_stride, c.ldin, P, st);
    else rc = split_planes_launch(feat, M, c.idim, c.ldin, fP, st);
    if (rc) return rc;
    GemmEpilogue ee = {c.embed_b, 0, 1.f, nullptr, 0, 0.f};
    if ((rc = tc_gemm(fP, m->embed_w, P, M, d, c.ldin, ee, y, d, nullptr, st))) return rc;
    if ((rc = layernorm_launch(y, c.embed_ln_g, c.embed_ln_b, x, M, d, 1e-5f, 1, st))) return rc;
    for (int l = 0; l < c.num_blocks; ++l) {
        const css_conformer_block& b = m->blk[l];
        const BlockPlanes& q = m->bp[l];
        bool att_ln_done = false;
        if ((rc = ffn_tc(b.ff_in, q.ff_in_w1, q.ff_in_w2, x, yP, hP, P, M, d, dffn, l > 0, b.att_ln_g, b.att_ln_b, &att_ln_done, st))) return rc;
        if (!att_ln_done &&
            (rc = layernorm_planes_launch(x, b.att_ln_g, b.att_ln_b, nullptr, M, d, 1e-5f, 0, yP.ptr, yP.plane_stride, P, st))) return rc;
        GemmEpilogue eq = {b.bqkv, 0, 1.f, nullptr, 0, 0.f};
        if (attention_tc_supported(T, d / c.heads)) {
            // Q, K, V leave the GEMM epilogue as bf16 planes (in the `h` region) and feed the tensor-core attention directly
            PlaneMat qP = {h, M * (long)(3 * d), 3 * d, P};
            if ((rc = tc_gemm(yP, q.wqkv, P, M, 3 * d, d, eq, nullptr, 0, &qP, st))) return rc;
            if ((rc = attention_tc_launch(qP.ptr, qP.plane_stride, P, c.pe ? m->pe.ptr : nullptr, m->pe.plane_stride, c.pe_rows,
                                          nullptr, aP.ptr, aP.plane_stride, P, N, T, c.heads, d / c.heads, st))) return rc;
        } else {
            if ((rc = tc_gemm(yP, q.wqkv, P, M, 3 * d, d, eq, h, 3 * d, nullptr, st))) return rc;
            if ((rc = attention_launch(h, c.pe, c.pe_rows, a, N, T, c.heads, d / c.heads, st))) return rc;
            if ((rc = split_planes_launch(a, M, d, d, aP, st))) return rc;
        }
        GemmEpilogue eo = {b.bo, 0, 1.f, x, d, 0.f};
        if ((rc = tc_gemm(aP, q.wo, P, M, d, d, eo, x, d, nullptr, st))) return rc;
        ConvModuleParams cp = {b.conv_ln_g, b.conv_ln_b, b.pw1_w0, b.pw1_b0, b.pw1_w1, b.pw1_b1, b.dw_w, b.dw_b,
                               b.bn_scale, b.bn_shift, b.pw2_w, b.pw2_b, 33};
        cp.dw_wt = m->taps_t + (size_t)l * 33 * d;
        if ((rc = conv_module_launch(x, a, N, T, d, cp, st))) return rc;
        { float* t = x; x = a; a = t; }
        // block output LN: fp32 for the residual stream, and with it the planes the next GEMM reads -- the next block's FFN-in
        // norm of that output, or the output itself for the head GEMM after the last block.  Both norms ride in the epilogue
        // of the FFN's second GEMM when the shape allows it (css_gemm_tc_ln_fusion(0) / CSS_B200_GEMM_LN=0: separate launch).
        const bool last = l == c.num_blocks - 1;
        const css_ffn_weights* nf = last ? nullptr : &m->blk[l + 1].ff_in;
        bool out_ln_done = false;
        if ((rc = ffn_tc(b.ff_out, q.ff_out_w1, q.ff_out_w2, x, yP, hP, P, M, d, dffn, false, b.out_ln_g,
                         b.out_ln_b, &out_ln_done, st, true, nf ? nf->ln_g : nullptr, nf ? nf->ln_b : nullptr))) return rc;
        if (!out_ln_done && (rc = layernorm2_planes_launch(x, b.out_ln_g, b.out_ln_b, x, nf ? nf->ln_g : nullptr, nf ? nf->ln_b : nullptr, M, d, 1e-5f,
                                           0, yP.ptr, yP.plane_stride, P, st)))
            return rc;
    }
    if (c.num_blocks == 0 && (rc = split_planes_launch(x, M, d, d, yP, st))) return rc;
    if (head) {
        GemmEpilogue eh = {c.head_b, 1, 1.f, nullptr, 0, 0.f};
        return tc_gemm(yP, m->head_w, P, M, nh, d, eh, head, ld_head, nullptr, st);
    }
    const int ldh = head_pitch(c);
    GemmEpilogue eh = {c.head_b, 1, 1.f, nullptr, 0, clamp_max};
    if ((rc = tc_gemm(yP, m->head_w, P, M, nh, d, eh, hd, ldh, nullptr, st))) return rc;
    return masks_layout_launch(hd, ldh, masks, N, T, c.num_bins, c.num_out, st);
}

extern "C" int css_conformer_forward(void* model, float* feat, long N, int T, int apply_mvn, float clamp_max,
                                     float* masks, float* head, int ld_head, void* workspace, size_t workspace_bytes,
                                     cudaStream_t st) {
    ConformerModel* m = static_cast<ConformerModel*>(model);
    CSS_REQUIRE(m && feat && workspace, "css_conformer_forward: null argument");
    CSS_REQUIRE(N > 0 && T > 1, "css_conformer_forward: empty input (N=%ld, T=%d)", N, T);
    CSS_REQUIRE((masks != nullptr) != (head != nullptr), "css_conformer_forward: pass exactly one of masks/head");
    CSS_REQUIRE(workspace_bytes >= css_conformer_workspace_bytes(model, N, T), "css_conformer_forward: workspace too small");
    const css_conformer_desc& c = m->d;
    const long M = N * T;
    const int d = c.d_model, dffn = c.d_ffn, nh = c.num_bins * c.num_out;
    const size_t wide = (size_t)(dffn > 3 * d ? dffn : 3 * d);
    char* wp = static_cast<char*>(workspace);
    if (g_gemm_mode != 0) return forward_tc(m, feat, nullptr, N, T, apply_mvn, clamp_max, masks, head, ld_head, wp, g_gemm_mode, st);
    float* x = (float*)wp;  wp += align256((size_t)M * d * 4);
    float* y = (float*)wp;  wp += align256((size_t)M * d * 4);
    float* a = (float*)wp;  wp += align256((size_t)M * d * 4);
    float* h = (float*)wp;  wp += align256((size_t)M * wide * 4);
    float* hd = (float*)wp;
    int rc;
    if (apply_mvn && (rc = css_mvn_time(feat, N, T, c.idim, c.ldin, st))) return rc;
    // embed: Linear -> LayerNorm -> ReLU   (css/models/conformer.py:124-129)
    GemmEpilogue ee = {c.embed_b, 0, 1.f, nullptr, 0, 0.f};
    if ((rc = gemm_dispatch(feat, c.ldin, c.embed_w, c.ldin, y, d, M, d, c.ldin, ee, st))) return rc;
    if ((rc = layernorm_launch(y, c.embed_ln_g, c.embed_ln_b, x, M, d, 1e-5f, 1, st))) return rc;
    for (int l = 0; l < c.num_blocks; ++l) {
        const css_conformer_block& b = m->blk[l];
        if ((rc = ffn(b.ff_in, x, y, h, M, d, dffn, st))) return rc;
        // x <- x + Wo attn(LN(x))     (css/models/conformer.py:213-254,351)
        if ((rc = layernorm_launch(x, b.att_ln_g, b.att_ln_b, y, M, d, 1e-5f, 0, st))) return rc;
        GemmEpilogue eq = {b.bqkv, 0, 1.f, nullptr, 0, 0.f};
        if ((rc = gemm_dispatch(y, d, b.wqkv, d, h, 3 * d, M, 3 * d, d, eq, st))) return rc;
        if ((rc = attention_launch(h, c.pe, c.pe_rows, a, N, T, c.heads, d / c.heads, st))) return rc;
        GemmEpilogue eo = {b.bo, 0, 1.f, x, d, 0.f};
        if ((rc = gemm_dispatch(a, d, b.wo, d, x, d, M, d, d, eo, st))) return rc;
        // x <- x + Conv(x)            (css/models/conformer.py:289-303,352) -- ping-pong x <-> a
        ConvModuleParams cp = {b.conv_ln_g, b.conv_ln_b, b.pw1_w0, b.pw1_b0, b.pw1_w1, b.pw1_b1, b.dw_w, b.dw_b,
                               b.bn_scale, b.bn_shift, b.pw2_w, b.pw2_b, 33};
        if ((rc = conv_module_launch(x, a, N, T, d, cp, st))) return rc;
        { float* t = x; x = a; a = t; }
        if ((rc = ffn(b.ff_out, x, y, h, M, d, dffn, st))) return rc;
        if ((rc = layernorm_launch(x, b.out_ln_g, b.out_ln_b, x, M, d, 1e-5f, 0, st))) return rc;
    }
    // head: Linear -> ReLU (-> clamp) -> 3 masks    (css/models/conformer.py:98-101, separator.py:100-104)
    if (head) {
        GemmEpilogue eh = {c.head_b, 1, 1.f, nullptr, 0, 0.f};
        return gemm_dispatch(x, d, c.head_w, d, head, ld_head, M, nh, d, eh, st);
    }
    const int ldh = head_pitch(c);
    GemmEpilogue eh = {c.head_b, 1, 1.f, nullptr, 0, clamp_max};
    if ((rc = gemm_dispatch(x, d, c.head_w, d, hd, ldh, M, nh, d, eh, st))) return rc;
    return masks_layout_launch(hd, ldh, masks, N, T, c.num_bins, c.num_out, st);
}

// The executor's path in one call: chunk spectra -> features -> input MVN -> network -> masks.  spec element
// (channel d, chunk b, frame t, bin k) lives at spec[(d*stride_d + b*stride_chunk + t) * 264 + k] (complex64), as css_features
// takes it; mag (N, 257, T) is the second output of css_features.  Tensor-core GEMM modes only (the fp32 SIMT mode keeps
// the two-call form css_features + css_conformer_forward).
extern "C" int css_conformer_forward_spec(void* model, const void* spec, long stride_d, long stride_chunk, const int* ipd_l,
                                          const int* ipd_r, int n_pairs, int shared_right, long N, int T, float clamp_max, float* mag, float* masks,
                                          float* head, int ld_head, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    ConformerModel* m = static_cast<ConformerModel*>(model);
    CSS_REQUIRE(m && spec && mag && workspace, "css_conformer_forward_spec: null argument");
    CSS_REQUIRE(N > 0 && T > 1, "css_conformer_forward_spec: empty input (N=%ld, T=%d)", N, T);
    CSS_REQUIRE((masks != nullptr) != (head != nullptr), "css_conformer_forward_spec: pass exactly one of masks/head");
    CSS_REQUIRE(g_gemm_mode != 0, "css_conformer_forward_spec: needs a tensor-core GEMM mode (css_set_gemm_mode 1 or 2)");
    CSS_REQUIRE(m->d.idim == CSS_NBINS * (n_pairs + 1), "css_conformer_forward_spec: model expects %d input features, %d IPD pairs give %d",
                m->d.idim, n_pairs, CSS_NBINS * (n_pairs + 1));
    CSS_REQUIRE(features_planes_supported(T, n_pairs), "css_conformer_forward_spec: T=%d with %d IPD pairs unsupported", T, n_pairs);
    CSS_REQUIRE(workspace_bytes >= css_conformer_workspace_bytes(model, N, T), "css_conformer_forward_spec: workspace too small");
    const SpecInput si = {spec, stride_d, stride_chunk, ipd_l, ipd_r, n_pairs, shared_right, mag};
    return forward_tc(m, nullptr, &si, N, T, 1, clamp_max, masks, head, ld_head, static_cast<char*>(workspace), g_gemm_mode, st);
}

extern "C" int css_features_planes_supported(int T, int n_pairs) { return features_planes_supported(T, n_pairs) ? 1 : 0; }
