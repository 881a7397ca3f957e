// Training-mode forward and backward of the Conformer mask estimator (css/models/conformer.py:83-104,152-165,348-362;
// step body css/trainer/iteration_types.py:34-63), orchestrated natively over flat parameter / gradient buffers.
// Every Linear (forward, dX = dY.W and dW = dY^T.X) runs on the tcgen05 GEMM over bf16 split planes; dW uses split-K
// with fp32 atomics because its output is small (<= 1024 x 1856) while its contraction runs over all N*T rows.
#include "common.cuh"
#include "conformer.h"
#include "gemm.h"
#include "train.h"
#include "../../include/css_b200.h"
#include <algorithm>
#include <new>
#include <vector>

namespace {

constexpr int kP = 2;   // planes kept for every operand (the GEMM mode decides how many are used)

struct LinPlanes { PlaneMat w; };   // W (N,K) planes [N][Kp]; the gradient GEMMs read them MN-major, no W^T copy
struct BlockLin { LinPlanes ff_in_w1, ff_in_w2, wqkv, wo, ff_out_w1, ff_out_w2; };

struct Trainer {
    css_conformer_train_desc d;
    std::vector<css_conformer_block_offsets> blk;
    float *params, *grads, *bn_running;
    void* plane_mem = nullptr;
    LinPlanes embed, head;
    std::vector<BlockLin> bl;
    int ldin, ldh, nhp;   // feature pitch of the planes (idim -> multiple of 64), fp32 head pitch, head K pitch (multiple of 64)
    float drop_p = 0.f;   // dropout probability of the six nn.Dropout sites (conformer.py:127,249,254,302,317,319)
    uint32_t drop_seed = 0;
    CssDrop drop(int block, int k) const { return css_make_drop(drop_p, drop_seed, css_drop_site(block, k)); }
};

inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }
inline long pad64(long v) { return (v + 63) / 64 * 64; }

struct SavedBlock {
    float *x0, *x1, *x2, *x3, *x4, *qkv, *cpre, *bnsave, *probs;
    PlaneMat n1, h1, n2, att, n4, h4;
};
struct Ws {
    long M, Mp;
    // saved by the forward
    PlaneMat fP, yP;
    float *ye, *xe, *xout, *head;
    std::vector<SavedBlock> sb;
    // scratch
    double* bn_stats;
    float* bstats;
    float *g, *gm, *dwide, *dn, *datt, *att_scratch, *tmp;
    PlaneMat gP, dwP;
    size_t bytes;
};

// carve the workspace; with base == nullptr only the size is computed
void carve(const Trainer* t, long N, int T, char* base, Ws* w) {
    const css_conformer_train_desc& c = t->d;
    const long M = N * T, Mp = pad64(M);
    const int d = c.d_model, f = c.d_ffn;
    const int wide = std::max(std::max(f, 3 * d), t->nhp);
    size_t off = 0;
    auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align256(bytes); return p; };
    auto planes = [&](long rows, int ld) { PlaneMat pm = {take((size_t)kP * rows * ld * 2), rows * (long)ld, ld, kP}; return pm; };
    auto f32 = [&](size_t n) { return reinterpret_cast<float*>(take(n * 4)); };
    w->M = M; w->Mp = Mp;
    w->fP = planes(M, t->ldin);
    w->yP = planes(M, d);
    w->ye = f32((size_t)M * d);
    w->xe = f32((size_t)M * d);
    w->xout = f32((size_t)M * d);
    w->head = f32((size_t)M * t->ldh);
    w->sb.resize(c.num_blocks);
    for (int l = 0; l < c.num_blocks; ++l) {
        SavedBlock& s = w->sb[l];
        s.x0 = l == 0 ? w->xe : f32((size_t)M * d);   // block input = LayerNorm output of the previous block
        s.x1 = f32((size_t)M * d); s.x2 = f32((size_t)M * d); s.x3 = f32((size_t)M * d); s.x4 = f32((size_t)M * d);
        s.qkv = f32((size_t)M * 3 * d);
        s.cpre = f32((size_t)M * d);
        s.bnsave = f32(2 * d);
        s.probs = f32(attention_train_probs_floats(N, T, c.heads));
        s.n1 = planes(M, d); s.h1 = planes(M, f); s.n2 = planes(M, d); s.att = planes(M, d); s.n4 = planes(M, d); s.h4 = planes(M, f);
    }
    w->bn_stats = reinterpret_cast<double*>(take(2 * d * sizeof(double)));
    w->bstats = f32(2 * d);
    w->g = f32((size_t)M * d);
    w->gm = f32((size_t)M * d);
    w->dwide = f32((size_t)M * wide);
    w->dn = f32((size_t)M * d);
    w->datt = f32((size_t)M * d);
    w->tmp = f32((size_t)M * d);
    w->att_scratch = f32(attention_train_scratch_floats(N, T, c.heads));
    w->gP = planes(M, d);
    w->dwP = planes(M, wide);
    w->bytes = off;
}

int tc(const PlaneMat& A, const PlaneMat& W, int P, long M, int N, int K, GemmEpilogue e, float* C, int ldc, const PlaneMat* out,
       cudaStream_t st) {
    css_prof_begin(0, 2.0 * (double)M * N * K, st);
    int rc = gemm_tc_planes(A, W, P, M, N, K, e, C, ldc, out, st);
    css_prof_end(st);
    return rc;
}

// view of the first `rows` x `ld` region of a plane buffer with a different logical shape
PlaneMat view(const PlaneMat& buf, long rows, int ld) {
    PlaneMat v = {buf.ptr, rows * (long)ld, ld, buf.planes};
    return v;
}

int build_planes(Trainer* t, cudaStream_t st) {
    const css_conformer_train_desc& c = t->d;
    const int d = c.d_model, f = c.d_ffn, nh = c.num_bins * c.num_out;
    int rc = 0;
    auto mk = [&](LinPlanes& lp, long off, int N, int K, bool need_T) {
        if (rc) return;
        (void)need_T;
        rc = split_t_launch(t->params + off, N, K, K, &lp.w, nullptr, 0, nullptr, 0.f, st);
    };
    mk(t->embed, c.embed_w, d, c.idim, false);
    mk(t->head, c.head_w, nh, d, true);
    for (int l = 0; l < c.num_blocks; ++l) {
        const css_conformer_block_offsets& o = t->blk[l];
        BlockLin& b = t->bl[l];
        mk(b.ff_in_w1, o.ff_in.w1, f, d, true);
        mk(b.ff_in_w2, o.ff_in.w2, d, f, true);
        mk(b.wqkv, o.wqkv, 3 * d, d, true);
        mk(b.wo, o.wo, d, d, true);
        mk(b.ff_out_w1, o.ff_out.w1, f, d, true);
        mk(b.ff_out_w2, o.ff_out.w2, d, f, true);
    }
    return rc;
}

int mode_planes() { return css_get_gemm_mode() == 2 ? 1 : 2; }

// The gradient GEMMs read their operands as they are stored (tcgen05 MN-major smem descriptors, GemmEpilogue::a_mn / b_mn):
//   dX = dY . W      : W (N_out, K_in) planes are the (K, N) matrix of that product          -> no W^T planes
//   dW += dY^T . X   : dY (M, N_out) and X (M, K_in) planes are its (K, M) and (K, N) matrices -> no transposed activations
// (the first version transposed every dY, X and W: 17.6 % of the step's kernel time)
GemmEpilogue ep_dx(float alpha, long k_valid) {
    GemmEpilogue e = {nullptr, 0, alpha, nullptr, 0, 0.f};
    e.b_mn = 1;
    e.k_valid = k_valid;
    return e;
}
GemmEpilogue ep_dw(float alpha, long rows) {
    GemmEpilogue e = {nullptr, 0, alpha, nullptr, 0, 0.f};
    e.a_mn = e.b_mn = 1;
    e.accumulate = 1;
    e.k_valid = rows;
    return e;
}

// blk / site: block index and the site id of the hidden-activation dropout (the output dropout is site + 1)
int ffn_fwd(Trainer* t, const css_ffn_offsets& o, const LinPlanes& w1, const LinPlanes& w2, const float* xin, float* xout,
            const PlaneMat& nP, const PlaneMat& hP, int P, long M, int blk, int site, float* tmp, cudaStream_t st) {
    const int d = t->d.d_model, f = t->d.d_ffn;
    const float* p = t->params;
    int rc = layernorm_planes_launch(xin, p + o.ln_g, p + o.ln_b, nullptr, M, d, 1e-5f, 0, nP.ptr, nP.plane_stride, kP, st);
    if (rc) return rc;
    GemmEpilogue e1 = {p + o.b1, 1, 1.f, nullptr, 0, 0.f};
    if ((rc = tc(nP, w1.w, P, M, f, d, e1, nullptr, 0, &hP, st))) return rc;
    if (t->drop_p <= 0.f) {
        GemmEpilogue e2 = {p + o.b2, 0, 0.5f, xin, d, 0.f};
        return tc(hP, w2.w, P, M, d, f, e2, xout, d, nullptr, st);
    }
    // net = Linear, ReLU, Dropout, Linear, Dropout (conformer.py:314-320): both dropouts as separate passes
    if ((rc = dropout_planes_launch(hP.ptr, hP.plane_stride, kP, M, f, hP.ld, t->drop(blk, site), st))) return rc;
    GemmEpilogue e2 = {p + o.b2, 0, 1.f, nullptr, 0, 0.f};
    if ((rc = tc(hP, w2.w, P, M, d, f, e2, tmp, d, nullptr, st))) return rc;
    return dropout_f32_launch(xout, tmp, xin, M * d, t->drop(blk, site + 1), 0.5f, 0, st);
}

// gradient of one macaron FFN: g (d loss / d x_out) -> g (d loss / d x_in), parameter gradients accumulated
int ffn_bwd(Trainer* t, const css_ffn_offsets& o, const LinPlanes& w1, const LinPlanes& w2, const float* xin, const PlaneMat& nP,
            const PlaneMat& hP, Ws& w, int P, int blk, int site, cudaStream_t st) {
    const int d = t->d.d_model, f = t->d.d_ffn;
    const long M = w.M, Mp = w.Mp;
    float* gr = t->grads;
    const float* p = t->params;
    int rc;
    PlaneMat dhP = view(w.dwP, M, f);
    const bool dr = t->drop_p > 0.f;
    const float hs = dr ? 1.f / (1.f - t->drop_p) : 1.f;          // hidden dropout: dh *= 1/(1-p) where kept (dropped h == 0 is masked)
    const float* gb = w.g;                                        // gradient w.r.t. the branch output W2 h + b2 (before the 0.5)
    if (dr) {
        if ((rc = dropout_f32_launch(w.gm, w.g, nullptr, M * d, t->drop(blk, site + 1), 1.f, 0, st))) return rc;
        gb = w.gm;
    }
    if ((rc = split_t_launch(gb, M, d, d, &w.gP, nullptr, 0, gr + o.b2, 0.5f, st))) return rc;
    // dW2 += 0.5 g^T h
    if ((rc = tc(w.gP, hP, P, d, f, (int)Mp, ep_dw(0.5f, M), gr + o.w2, f, nullptr, st))) return rc;
    // dh = 0.5 g W2, masked by the ReLU
    GemmEpilogue eh = ep_dx(0.5f * hs, d);
    eh.relu_mask = hP.ptr;
    eh.ldmask = hP.ld;
    if ((rc = tc(w.gP, w2.w, P, M, f, d, eh, w.dwide, f, nullptr, st))) return rc;
    if ((rc = split_t_launch(w.dwide, M, f, f, &dhP, nullptr, 0, gr + o.b1, 1.f, st))) return rc;
    // dW1 += dh^T n
    if ((rc = tc(dhP, nP, P, f, d, (int)Mp, ep_dw(1.f, M), gr + o.w1, d, nullptr, st))) return rc;
    // dn = dh W1 ; g += LN'(x_in) dn
    if ((rc = tc(dhP, w1.w, P, M, d, f, ep_dx(1.f, f), w.dn, d, nullptr, st))) return rc;
    LnBwdArgs a = {xin, p + o.ln_g, nullptr, w.dn, nullptr, w.g, 1, gr + o.ln_g, gr + o.ln_b, nullptr, nullptr, nullptr, nullptr};
    return ln_bwd_launch(a, M, d, 1e-5f, false, st);
}

ConvTrainParams conv_params(Trainer* t, int l) {
    const css_conformer_block_offsets& o = t->blk[l];
    const float* p = t->params;
    float* rm = t->bn_running ? t->bn_running + (size_t)l * 2 * t->d.d_model : nullptr;
    ConvTrainParams cp = {p + o.conv_ln_g, p + o.conv_ln_b, p + o.pw1_w, p + o.pw1_b, p + o.dw_w, p + o.dw_b, p + o.bn_g,
                          p + o.bn_b, p + o.pw2_w, p + o.pw2_b, rm, rm ? rm + t->d.d_model : nullptr};
    return cp;
}
ConvTrainGrads conv_grads(Trainer* t, int l) {
    const css_conformer_block_offsets& o = t->blk[l];
    float* g = t->grads;
    ConvTrainGrads cg = {g + o.conv_ln_g, g + o.conv_ln_b, g + o.pw1_w, g + o.pw1_b, g + o.dw_w, g + o.dw_b, g + o.bn_g,
                         g + o.bn_b, g + o.pw2_w, g + o.pw2_b};
    return cg;
}

}  // namespace

extern "C" void* css_trainer_create(const css_conformer_train_desc* desc, const css_conformer_block_offsets* blocks, float* params,
                                    float* grads, float* bn_running) {
    if (!desc || !blocks || !params || !grads) { css_set_error("css_trainer_create: null argument"); return nullptr; }
    if (desc->d_model != 256 || desc->d_model / desc->heads != 64 || desc->ksize != 33 || desc->d_ffn % 128 || desc->d_ffn > 1024) {
        css_set_error("css_trainer_create: kernels are built for attention_dim 256, 64-wide heads, kernel_size 33, linear_units "
                      "<= 1024 (got d=%d heads=%d ksize=%d ffn=%d)", desc->d_model, desc->heads, desc->ksize, desc->d_ffn);
        return nullptr;
    }
    Trainer* t = new (std::nothrow) Trainer();
    if (!t) { css_set_error("css_trainer_create: out of host memory"); return nullptr; }
    t->d = *desc;
    t->blk.assign(blocks, blocks + desc->num_blocks);
    t->params = params; t->grads = grads; t->bn_running = bn_running;
    const int d = desc->d_model, f = desc->d_ffn, nh = desc->num_bins * desc->num_out;
    t->ldin = (int)pad64(desc->idim);
    t->ldh = (nh + 7) / 8 * 8;
    t->nhp = (int)pad64(nh);
    // plane memory: W planes [N][pad64(K)] and W^T planes [K][pad64(N)] of every Linear
    size_t total = 0;
    auto bytes = [](long rows, long ld) { return align256((size_t)kP * rows * ld * 2); };
    auto size_lin = [&](int N, int K, bool) { total += bytes(N, pad64(K)); };
    size_lin(d, desc->idim, false);
    size_lin(nh, d, true);
    for (int l = 0; l < desc->num_blocks; ++l) {
        size_lin(f, d, true); size_lin(d, f, true); size_lin(3 * d, d, true); size_lin(d, d, true); size_lin(f, d, true); size_lin(d, f, true);
    }
    cudaError_t e = cudaMalloc(&t->plane_mem, total);
    if (e != cudaSuccess) {
        css_set_error("css_trainer_create: cannot allocate %zu B of weight planes: %s", total, cudaGetErrorString(e));
        delete t;
        return nullptr;
    }
    char* wp = static_cast<char*>(t->plane_mem);
    auto mk = [&](int N, int K, bool T) {
        LinPlanes lp{};
        const long kp = pad64(K), np = pad64(N);
        lp.w = {wp, N * kp, (int)kp, kP};
        wp += bytes(N, kp);
        (void)T; (void)np;
        return lp;
    };
    t->embed = mk(d, desc->idim, false);
    t->head = mk(nh, d, true);
    t->bl.resize(desc->num_blocks);
    for (int l = 0; l < desc->num_blocks; ++l) {
        BlockLin& b = t->bl[l];
        b.ff_in_w1 = mk(f, d, true); b.ff_in_w2 = mk(d, f, true); b.wqkv = mk(3 * d, d, true); b.wo = mk(d, d, true);
        b.ff_out_w1 = mk(f, d, true); b.ff_out_w2 = mk(d, f, true);
    }
    return t;
}

extern "C" int css_trainer_set_dropout(void* trainer, float p, unsigned seed) {
    Trainer* t = static_cast<Trainer*>(trainer);
    CSS_REQUIRE(t && p >= 0.f && p < 1.f, "css_trainer_set_dropout: probability %g out of [0, 1)", (double)p);
    t->drop_p = p;
    t->drop_seed = seed;
    return 0;
}

extern "C" void css_trainer_destroy(void* trainer) {
    Trainer* t = static_cast<Trainer*>(trainer);
    if (t && t->plane_mem) cudaFree(t->plane_mem);
    delete t;
}

extern "C" size_t css_trainer_workspace_bytes(const void* trainer, long N, int T) {
    Ws w;
    carve(static_cast<const Trainer*>(trainer), N, T, nullptr, &w);
    return w.bytes;
}

extern "C" int css_trainer_forward(void* trainer, const float* feat, long N, int T, float* y_pred, float* masks_head, int ld_head,
                                   void* workspace, size_t workspace_bytes, cudaStream_t st) {
    Trainer* t = static_cast<Trainer*>(trainer);
    CSS_REQUIRE(t && feat && y_pred && workspace, "css_trainer_forward: null argument");
    CSS_REQUIRE(N > 0 && T > 1 && T <= 256, "css_trainer_forward: bad shape N=%ld T=%d (T <= 256)", N, T);
    CSS_REQUIRE(css_get_gemm_mode() != 0, "css_trainer_forward: the training path runs on the tcgen05 GEMM (gemm mode 1 or 2)");
    Ws w;
    carve(t, N, T, static_cast<char*>(workspace), &w);
    CSS_REQUIRE(workspace_bytes >= w.bytes, "css_trainer_forward: workspace too small (%zu < %zu)", workspace_bytes, w.bytes);
    const css_conformer_train_desc& c = t->d;
    const float* p = t->params;
    const long M = w.M;
    const int d = c.d_model, nh = c.num_bins * c.num_out, P = mode_planes();
    int rc;
    if ((rc = build_planes(t, st))) return rc;
    if ((rc = mvn_time_planes_launch(feat, N, T, c.idim, c.idim, w.fP.ptr, w.fP.plane_stride, t->ldin, kP, st))) return rc;
    GemmEpilogue ee = {p + c.embed_b, 0, 1.f, nullptr, 0, 0.f};
    if ((rc = tc(w.fP, t->embed.w, P, M, d, t->ldin, ee, w.ye, d, nullptr, st))) return rc;
    const bool dr = t->drop_p > 0.f;
    if (!dr) {
        if ((rc = layernorm_launch(w.ye, p + c.embed_ln_g, p + c.embed_ln_b, w.xe, M, d, 1e-5f, 1, st))) return rc;
    } else {                                      // embed = Linear, LayerNorm, Dropout, ReLU (conformer.py:124-128)
        if ((rc = layernorm_launch(w.ye, p + c.embed_ln_g, p + c.embed_ln_b, w.tmp, M, d, 1e-5f, 0, st))) return rc;
        if ((rc = dropout_f32_launch(w.xe, w.tmp, nullptr, M * d, t->drop(-1, 0), 1.f, 1, st))) return rc;
    }
    const float* pe = c.pe >= 0 ? p + c.pe : nullptr;
    for (int l = 0; l < c.num_blocks; ++l) {
        const css_conformer_block_offsets& o = t->blk[l];
        const BlockLin& b = t->bl[l];
        SavedBlock& s = w.sb[l];
        if ((rc = ffn_fwd(t, o.ff_in, b.ff_in_w1, b.ff_in_w2, s.x0, s.x1, s.n1, s.h1, P, M, l, 0, w.dn, st))) return rc;
        if ((rc = layernorm_planes_launch(s.x1, p + o.att_ln_g, p + o.att_ln_b, nullptr, M, d, 1e-5f, 0, s.n2.ptr, s.n2.plane_stride, kP, st))) return rc;
        GemmEpilogue eq = {p + o.bqkv, 0, 1.f, nullptr, 0, 0.f};
        if ((rc = tc(s.n2, b.wqkv.w, P, M, 3 * d, d, eq, s.qkv, 3 * d, nullptr, st))) return rc;
        const CssDrop dp = t->drop(l, 2), dc = t->drop(l, 4);
        if ((rc = attention_train_fwd_launch(s.qkv, pe, c.pe_rows, w.tmp, s.probs, w.att_scratch, N, T, c.heads, d / c.heads, P == 2,
                                             dr ? &dp : nullptr, st))) return rc;
        if ((rc = split_planes_launch(w.tmp, M, d, d, s.att, st))) return rc;
        if (!dr) {
            GemmEpilogue eo = {p + o.bo, 0, 1.f, s.x1, d, 0.f};
            if ((rc = tc(s.att, b.wo.w, P, M, d, d, eo, s.x2, d, nullptr, st))) return rc;
        } else {                                  // x + dropout(linear_out(.)) (conformer.py:254)
            GemmEpilogue eo = {p + o.bo, 0, 1.f, nullptr, 0, 0.f};
            if ((rc = tc(s.att, b.wo.w, P, M, d, d, eo, w.tmp, d, nullptr, st))) return rc;
            if ((rc = dropout_f32_launch(s.x2, w.tmp, s.x1, M * d, t->drop(l, 3), 1.f, 0, st))) return rc;
        }
        if ((rc = conv_train_fwd_launch(s.x2, s.x3, s.cpre, s.bnsave, w.bn_stats, N, T, conv_params(t, l), dr ? &dc : nullptr, st))) return rc;
        if ((rc = ffn_fwd(t, o.ff_out, b.ff_out_w1, b.ff_out_w2, s.x3, s.x4, s.n4, s.h4, P, M, l, 5, w.dn, st))) return rc;
        const bool last = l == c.num_blocks - 1;
        float* xo = last ? w.xout : w.sb[l + 1].x0;
        if ((rc = layernorm_planes_launch(s.x4, p + o.out_ln_g, p + o.out_ln_b, xo, M, d, 1e-5f, 0, last ? w.yP.ptr : nullptr,
                                          w.yP.plane_stride, last ? kP : 0, st))) return rc;
    }
    if (c.num_blocks == 0 && (rc = split_planes_launch(w.xe, M, d, d, w.yP, st))) return rc;
    GemmEpilogue eh = {p + c.head_b, 1, 1.f, nullptr, 0, 0.f};
    if ((rc = tc(w.yP, t->head.w, P, M, nh, d, eh, w.head, t->ldh, nullptr, st))) return rc;
    if (masks_head) {
        CSS_REQUIRE(ld_head >= nh, "css_trainer_forward: ld_head too small");
        cudaMemcpy2DAsync(masks_head, (size_t)ld_head * 4, w.head, (size_t)t->ldh * 4, (size_t)nh * 4, M, cudaMemcpyDeviceToDevice, st);
    }
    return ypred_launch(w.head, t->ldh, feat, c.idim, y_pred, N, T, c.num_bins, c.num_spk, st);
}

// parts: num_blocks = mask head, l = encoder block l, -1 = input embedding; processed from part_hi down to part_lo
extern "C" int css_trainer_backward_part(void* trainer, const float* feat, const float* gy, long N, int T, int part_hi, int part_lo,
                                         void* workspace, size_t workspace_bytes, cudaStream_t st) {
    Trainer* t = static_cast<Trainer*>(trainer);
    CSS_REQUIRE(t && feat && workspace, "css_trainer_backward: null argument");
    Ws w;
    carve(t, N, T, static_cast<char*>(workspace), &w);
    CSS_REQUIRE(workspace_bytes >= w.bytes, "css_trainer_backward: workspace too small");
    const css_conformer_train_desc& c = t->d;
    CSS_REQUIRE(part_hi <= c.num_blocks && part_lo >= -1 && part_lo <= part_hi, "css_trainer_backward: bad part range");
    const float* p = t->params;
    float* gr = t->grads;
    const long M = w.M, Mp = w.Mp;
    const int d = c.d_model, nh = c.num_bins * c.num_out, P = mode_planes();
    const float* pe = c.pe >= 0 ? p + c.pe : nullptr;
    int rc;
    for (int part = part_hi; part >= part_lo; --part) {
        if (part == c.num_blocks) {
            // ---- mask head: y_pred = relu(x W^T + b)[:, :num_spk*nb] * f_orig ----
            CSS_REQUIRE(gy, "css_trainer_backward: null gradient");
            float* dhead = w.dwide;   // (M, ldh)
            if ((rc = head_bwd_launch(w.head, t->ldh, feat, c.idim, gy, dhead, N, T, c.num_bins, c.num_spk, st))) return rc;
            PlaneMat dhP = view(w.dwP, M, t->nhp);
            if ((rc = split_t_launch(dhead, M, nh, t->ldh, &dhP, nullptr, 0, gr + c.head_b, 1.f, st))) return rc;
            if ((rc = tc(dhP, w.yP, P, nh, d, (int)Mp, ep_dw(1.f, M), gr + c.head_w, d, nullptr, st))) return rc;
            if ((rc = tc(dhP, t->head.w, P, M, d, t->nhp, ep_dx(1.f, nh), w.g, d, nullptr, st))) return rc;
        } else if (part >= 0) {
            const int l = part;
            const css_conformer_block_offsets& o = t->blk[l];
            const BlockLin& b = t->bl[l];
            SavedBlock& s = w.sb[l];
            // block output LayerNorm
            LnBwdArgs a0 = {s.x4, p + o.out_ln_g, nullptr, w.g, nullptr, w.g, 0, gr + o.out_ln_g, gr + o.out_ln_b, nullptr, nullptr, nullptr, nullptr};
            if ((rc = ln_bwd_launch(a0, M, d, 1e-5f, false, st))) return rc;
            const bool dr = t->drop_p > 0.f;
            if ((rc = ffn_bwd(t, o.ff_out, b.ff_out_w1, b.ff_out_w2, s.x3, s.n4, s.h4, w, P, l, 5, st))) return rc;
            const float* gconv = w.g;
            if (dr) {
                if ((rc = dropout_f32_launch(w.gm, w.g, nullptr, M * d, t->drop(l, 4), 1.f, 0, st))) return rc;
                gconv = w.gm;
            }
            if ((rc = conv_train_bwd_launch(s.x2, s.cpre, s.bnsave, w.g, gconv, w.dn, w.bstats, N, T, conv_params(t, l), conv_grads(t, l), st))) return rc;
            // ---- self-attention ----
            const float* gatt = w.g;
            if (dr) {
                if ((rc = dropout_f32_launch(w.gm, w.g, nullptr, M * d, t->drop(l, 3), 1.f, 0, st))) return rc;
                gatt = w.gm;
            }
            if ((rc = split_t_launch(gatt, M, d, d, &w.gP, nullptr, 0, gr + o.bo, 1.f, st))) return rc;
            if ((rc = tc(w.gP, s.att, P, d, d, (int)Mp, ep_dw(1.f, M), gr + o.wo, d, nullptr, st))) return rc;
            if ((rc = tc(w.gP, b.wo.w, P, M, d, d, ep_dx(1.f, d), w.datt, d, nullptr, st))) return rc;
            float* dqkv = w.dwide;   // (M, 3d)
            const CssDrop dp = t->drop(l, 2);
            if ((rc = attention_train_bwd_launch(s.qkv, pe, c.pe_rows, s.probs, w.datt, dqkv, pe ? gr + c.pe : nullptr, w.att_scratch,
                                                 N, T, c.heads, d / c.heads, P == 2, dr ? &dp : nullptr, st))) return rc;
            PlaneMat dqP = view(w.dwP, M, 3 * d);
            if ((rc = split_t_launch(dqkv, M, 3 * d, 3 * d, &dqP, nullptr, 0, gr + o.bqkv, 1.f, st))) return rc;
            if ((rc = tc(dqP, s.n2, P, 3 * d, d, (int)Mp, ep_dw(1.f, M), gr + o.wqkv, d, nullptr, st))) return rc;
            if ((rc = tc(dqP, b.wqkv.w, P, M, d, 3 * d, ep_dx(1.f, 3 * d), w.dn, d, nullptr, st))) return rc;
            LnBwdArgs a1 = {s.x1, p + o.att_ln_g, nullptr, w.dn, nullptr, w.g, 1, gr + o.att_ln_g, gr + o.att_ln_b, nullptr, nullptr, nullptr, nullptr};
            if ((rc = ln_bwd_launch(a1, M, d, 1e-5f, false, st))) return rc;
            if ((rc = ffn_bwd(t, o.ff_in, b.ff_in_w1, b.ff_in_w2, s.x0, s.n1, s.h1, w, P, l, 0, st))) return rc;
        } else {
            // ---- input embedding: x_e = relu(dropout(LN(f W^T + b))) ----
            if (t->drop_p > 0.f) {        // kept elements carry 1/(1-p); dropped ones have x_e == 0 and fall to the ReLU mask
                const CssDrop none = {0, 0, 1.f / (1.f - t->drop_p)};
                if ((rc = dropout_f32_launch(w.g, w.g, nullptr, M * d, none, 1.f, 0, st))) return rc;
            }
            LnBwdArgs a = {w.ye, p + c.embed_ln_g, nullptr, w.g, w.xe, w.g, 0, gr + c.embed_ln_g, gr + c.embed_ln_b, nullptr, nullptr, nullptr, nullptr};
            if ((rc = ln_bwd_launch(a, M, d, 1e-5f, false, st))) return rc;
            if ((rc = split_t_launch(w.g, M, d, d, &w.gP, nullptr, 0, gr + c.embed_b, 1.f, st))) return rc;
            if ((rc = tc(w.gP, w.fP, P, d, c.idim, (int)Mp, ep_dw(1.f, M), gr + c.embed_w, c.idim, nullptr, st))) return rc;
        }
    }
    return 0;
}

extern "C" int css_trainer_backward(void* trainer, const float* feat, const float* gy, long N, int T, void* workspace,
                                    size_t workspace_bytes, cudaStream_t st) {
    Trainer* t = static_cast<Trainer*>(trainer);
    CSS_REQUIRE(t, "css_trainer_backward: null trainer");
    return css_trainer_backward_part(trainer, feat, gy, N, T, t->d.num_blocks, -1, workspace, workspace_bytes, st);
}
