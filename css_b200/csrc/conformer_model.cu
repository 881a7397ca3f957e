// Conformer mask estimator forward (css/models/conformer.py:83-104,152-165,348-362), orchestrated
// natively: one C call launches the whole 16-block network on the caller's stream.
#include "common.cuh"
#include "conformer.h"
#include "gemm.h"
#include "../../include/css_b200.h"
#include <new>
#include <vector>

static int g_gemm_mode = 0;
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

struct ConformerModel {
    css_conformer_desc d;
    std::vector<css_conformer_block> blk;
};

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

extern "C" void css_conformer_destroy(void* model) { delete static_cast<ConformerModel*>(model); }

static inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }
static inline int head_pitch(const css_conformer_desc& d) { return ((d.num_bins * d.num_out + 7) / 8) * 8; }

extern "C" size_t css_conformer_workspace_bytes(const void* model, long N, int T) {
    const ConformerModel* m = static_cast<const ConformerModel*>(model);
    const size_t M = (size_t)N * T, d = m->d.d_model;
    const size_t wide = (size_t)(m->d.d_ffn > 3 * m->d.d_model ? m->d.d_ffn : 3 * m->d.d_model);
    return 3 * align256(M * d * 4) + align256(M * wide * 4) + align256(M * head_pitch(m->d) * 4);
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
