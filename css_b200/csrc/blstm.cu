// BLSTM mask estimator forward (css/models/blstm.py:68-88,122-128,156-160).
// Per layer: one GEMM projects all T inputs of both directions (x W_ih^T + b_ih + b_hh), then T
// recurrence steps each run ONE kernel for both directions: a (chunks x H) . (H x 4H) product
// tiled 64 x (32 hidden units x 4 gates) with the LSTM cell fused in the epilogue, so the gates
// never touch HBM.  Chunks are the batch axis that keeps the step kernel wide.
#include "common.cuh"
#include "conformer.h"
#include "gemm.h"
#include "../../include/css_b200.h"
#include <new>
#include <vector>

int gemm_dispatch(const float* A, int lda, const float* W, int ldw, float* C, int ldc, long M, int N, int K,
                  GemmEpilogue ep, cudaStream_t st);

#define LS_BN 64     // sequences per CTA
#define LS_BJ 32     // hidden units per CTA (x4 gates = 128 columns)
#define LS_BK 16

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// xproj: (N, T, 8H) = [fwd i f g o | bwd i f g o];  w_hh[dir]: (4H, H);  h_prev/h_next: (2, N, H);
// c: (2, N, H);  out: (N, T, 2H) = [fwd h | bwd h].   step s: fwd t = s, bwd t = T-1-s.
__global__ void __launch_bounds__(256)
lstm_step_kernel(const float* __restrict__ xproj, const float* __restrict__ w_hh_f, const float* __restrict__ w_hh_b,
                 const float* __restrict__ h_prev, float* __restrict__ h_next, float* __restrict__ c_state,
                 float* __restrict__ out, long N, int T, int H, int s) {
    __shared__ __align__(16) float Hs[2][LS_BK][LS_BN + 4];
    __shared__ __align__(16) float Ws[2][LS_BK][4 * LS_BJ + 4];
    const int dir = blockIdx.z;
    const int t = dir ? (T - 1 - s) : s;
    const float* W = dir ? w_hh_b : w_hh_f;
    const float* hp = h_prev + (long)dir * N * H;
    const long n0 = (long)blockIdx.x * LS_BN;
    const int j0 = blockIdx.y * LS_BJ;
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;       // thread: rows ty*4..+3, hidden units 2tx, 2tx+1
    const int lk = (tid & 3) * 4;
    const int hrow = tid >> 2;                    // 0..63 : sequence row for the H tile
    float4 ph, pw[2];
    auto gload = [&](int k0) {
        const long n = n0 + hrow;
        ph = (n < N) ? *reinterpret_cast<const float4*>(hp + n * H + k0 + lk) : make_float4(0, 0, 0, 0);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int col = hrow + 64 * i;        // 0..127 = gate*32 + jj
            const int row = (col >> 5) * H + j0 + (col & 31);
            pw[i] = *reinterpret_cast<const float4*>(W + (long)row * H + k0 + lk);
        }
    };
    auto sstore = [&](int buf) {
        Hs[buf][lk + 0][hrow] = ph.x; Hs[buf][lk + 1][hrow] = ph.y;
        Hs[buf][lk + 2][hrow] = ph.z; Hs[buf][lk + 3][hrow] = ph.w;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int col = hrow + 64 * i;
            Ws[buf][lk + 0][col] = pw[i].x; Ws[buf][lk + 1][col] = pw[i].y;
            Ws[buf][lk + 2][col] = pw[i].z; Ws[buf][lk + 3][col] = pw[i].w;
        }
    };
    float acc[4][4][2];   // [row][gate][unit]
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int g = 0; g < 4; ++g) acc[r][g][0] = acc[r][g][1] = 0.f;
    gload(0);
    sstore(0);
    __syncthreads();
    const int nk = H / LS_BK;
    for (int kt = 0; kt < nk; ++kt) {
        const int cur = kt & 1;
        if (kt + 1 < nk) gload((kt + 1) * LS_BK);
#pragma unroll
        for (int k = 0; k < LS_BK; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(&Hs[cur][k][ty * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const float2 w = *reinterpret_cast<const float2*>(&Ws[cur][k][g * LS_BJ + 2 * tx]);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    acc[r][g][0] = fmaf(av[r], w.x, acc[r][g][0]);
                    acc[r][g][1] = fmaf(av[r], w.y, acc[r][g][1]);
                }
            }
        }
        if (kt + 1 < nk) {
            sstore(cur ^ 1);
            __syncthreads();
        }
    }
    // fused cell update (gate order i, f, g, o as torch.nn.LSTM)
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const long n = n0 + ty * 4 + r;
        if (n >= N) continue;
        const float* xp = xproj + (n * T + t) * (long)(8 * H) + (long)dir * 4 * H;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int j = j0 + 2 * tx + u;
            const float gi = acc[r][0][u] + xp[j];
            const float gf = acc[r][1][u] + xp[H + j];
            const float gg = acc[r][2][u] + xp[2 * H + j];
            const float go = acc[r][3][u] + xp[3 * H + j];
            const long si = ((long)dir * N + n) * H + j;
            const float c = sigmoidf_(gf) * c_state[si] + sigmoidf_(gi) * tanhf(gg);
            const float h = sigmoidf_(go) * tanhf(c);
            c_state[si] = c;
            h_next[si] = h;
            out[(n * T + t) * (long)(2 * H) + (long)dir * H + j] = h;
        }
    }
}

struct BlstmModel {
    css_blstm_desc d;
    std::vector<css_blstm_layer> lay;
};

extern "C" void* css_blstm_create(const css_blstm_desc* desc, const css_blstm_layer* layers) {
    if (!desc || !layers) { css_set_error("css_blstm_create: null descriptor"); return nullptr; }
    if (desc->hdim % 128 || desc->hdim > 1024 || (desc->hdim / 2) % LS_BJ || desc->ldin % 16 || desc->ldin < desc->idim) {
        css_set_error("css_blstm_create: unsupported hdim=%d / ldin=%d", desc->hdim, desc->ldin);
        return nullptr;
    }
    BlstmModel* m = new (std::nothrow) BlstmModel();
    if (!m) { css_set_error("css_blstm_create: out of host memory"); return nullptr; }
    m->d = *desc;
    m->lay.assign(layers, layers + desc->num_layers);
    return m;
}
extern "C" void css_blstm_destroy(void* model) { delete static_cast<BlstmModel*>(model); }

static inline size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }
static inline int bl_head_pitch(const css_blstm_desc& d) { return ((d.num_bins * d.num_out + 7) / 8) * 8; }

extern "C" size_t css_blstm_workspace_bytes(const void* model, long N, int T) {
    const BlstmModel* m = static_cast<const BlstmModel*>(model);
    const size_t M = (size_t)N * T, h = m->d.hdim;
    return 2 * al256(M * h * 4) + al256(M * 4 * h * 4) + al256((size_t)N * h * 4) * 3 +
           al256(M * bl_head_pitch(m->d) * 4);
}

extern "C" int css_blstm_forward(void* model, float* feat, long N, int T, int apply_mvn, float clamp_max, float* masks,
                                 float* head, int ld_head, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    BlstmModel* m = static_cast<BlstmModel*>(model);
    CSS_REQUIRE(m && feat && workspace, "css_blstm_forward: null argument");
    CSS_REQUIRE(N > 0 && T > 1, "css_blstm_forward: empty input (N=%ld, T=%d)", N, T);
    CSS_REQUIRE((masks != nullptr) != (head != nullptr), "css_blstm_forward: pass exactly one of masks/head");
    CSS_REQUIRE(workspace_bytes >= css_blstm_workspace_bytes(model, N, T), "css_blstm_forward: workspace too small");
    const css_blstm_desc& c = m->d;
    const long M = N * T;
    const int h = c.hdim, H = h / 2, nh = c.num_bins * c.num_out;
    char* wp = static_cast<char*>(workspace);
    float* x = (float*)wp;   wp += al256((size_t)M * h * 4);
    float* y = (float*)wp;   wp += al256((size_t)M * h * 4);
    float* xp = (float*)wp;  wp += al256((size_t)M * 4 * h * 4);
    float* hs0 = (float*)wp; wp += al256((size_t)N * h * 4);
    float* hs1 = (float*)wp; wp += al256((size_t)N * h * 4);
    float* cs = (float*)wp;  wp += al256((size_t)N * h * 4);
    float* hd = (float*)wp;
    int rc;
    if (apply_mvn && (rc = css_mvn_time(feat, N, T, c.idim, c.ldin, st))) return rc;
    GemmEpilogue ee = {c.embed_b, 0, 1.f, nullptr, 0, 0.f};
    if ((rc = gemm_dispatch(feat, c.ldin, c.embed_w, c.ldin, y, h, M, h, c.ldin, ee, st))) return rc;
    if ((rc = layernorm_launch(y, c.embed_ln_g, c.embed_ln_b, x, M, h, 1e-5f, 1, st))) return rc;
    for (int l = 0; l < c.num_layers; ++l) {
        const css_blstm_layer& L = m->lay[l];
        GemmEpilogue ep = {L.b_gates, 0, 1.f, nullptr, 0, 0.f};
        if ((rc = gemm_dispatch(x, h, L.w_ih, h, xp, 4 * h, M, 4 * h, h, ep, st))) return rc;
        cudaMemsetAsync(hs0, 0, (size_t)N * h * 4, st);
        cudaMemsetAsync(cs, 0, (size_t)N * h * 4, st);
        dim3 grid((unsigned)((N + LS_BN - 1) / LS_BN), H / LS_BJ, 2);
        float *hp = hs0, *hn = hs1;
        for (int s = 0; s < T; ++s) {
            css_prof_begin(4, 2.0 * 2.0 * (double)N * H * 4 * H, st);
            lstm_step_kernel<<<grid, 256, 0, st>>>(xp, L.w_hh_f, L.w_hh_b, hp, hn, cs, y, N, T, H, s);
            css_prof_end(st);
            if ((rc = css_check_launch("lstm_step"))) return rc;
            float* tmp = hp; hp = hn; hn = tmp;
        }
        if ((rc = layernorm_launch(y, L.ln_g, L.ln_b, x, M, h, 1e-5f, 0, st))) return rc;
    }
    if (head) {
        GemmEpilogue eh = {c.head_b, 1, 1.f, nullptr, 0, 0.f};
        return gemm_dispatch(x, h, c.head_w, h, head, ld_head, M, nh, h, eh, st);
    }
    const int ldh = bl_head_pitch(c);
    GemmEpilogue eh = {c.head_b, 1, 1.f, nullptr, 0, clamp_max};
    if ((rc = gemm_dispatch(x, h, c.head_w, h, hd, ldh, M, nh, h, eh, st))) return rc;
    return masks_layout_launch(hd, ldh, masks, N, T, c.num_bins, c.num_out, st);
}
