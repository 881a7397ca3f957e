// Relative-position self-attention for the TRAINING step (css/models/conformer.py:229-250), forward and backward,
// for sequences up to 256 frames (training windows are 250 frames; the inference kernel in attention_tc.cu covers
// the executor's 150-frame chunks).
//
// Everything is expressed as strided-batched matrix products on the warp-level tensor cores (mma.sync m16n8k8 TF32,
// fp32 operands straight from memory; with SPLIT every product is the three-term split hi*hi + hi*lo + lo*hi, which
// keeps fp32-class accuracy) plus two row-wise kernels that own the relative-position skew:
//
//   forward   R = Q PE_win^T                (T x 2T-1)   PE_win[rho] = pe[rho - (T-1) + maxlen]
//             S = Q K^T                     (T x T)
//             P = softmax_j((S[i][j] + R[i][i-j+T-1]) / sqrt(dk))            -> kept for the backward
//             O = P V
//   backward  dP = dO V^T ; dS = P (dP - rowsum(P dP)) / sqrt(dk) ; dR[i][i-j+T-1] = dS[i][j]
//             dV = P^T dO ; dK = dS^T Q ; dQ = dS K + dR PE_win ; dPE_win = sum_(n,i) dR^T Q   (split-K, atomics)
#include "common.cuh"
#include "train.h"

namespace {

struct BMat {            // element (r, c) of batch (h, n): p[h*s_h + n*s_n + r*ld + c]
    const float* p;
    long s_h, s_n;
    int ld;
};
struct BOut {
    float* p;
    long s_h, s_n;
    int ld;
};

constexpr int BT = 64;       // block tile (M and N)
constexpr int BK = 32;
constexpr int PITCH = BK + 4;

__device__ __forceinline__ uint32_t to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void mma_tf32(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// C[b] (=, +=, atomic +=) alpha * op(A[b]) op(B[b])^T.   op(A) is M x K: stored [m][k] (TA = false) or [k][m] (TA = true);
// op(B) is N x K: stored [n][k] (TB = false) or [k][n] (TB = true).  grid = (n tiles, m tiles, batches * ksplit).
template <bool TA, bool TB, bool SPLIT>
__global__ void __launch_bounds__(128)
bgemm_kernel(BMat A, BMat B, BOut C, int M, int N, int K, int Nb, float alpha, int acc_mode, int ksplit) {
    __shared__ __align__(16) float As[BT][PITCH];
    __shared__ __align__(16) float Bs[BT][PITCH];
    const int split = blockIdx.z % ksplit, b = blockIdx.z / ksplit;
    const int h = b / Nb, n = b - h * Nb;
    const float* Ap = A.p + h * A.s_h + n * A.s_n;
    const float* Bp = B.p + h * B.s_h + n * B.s_n;
    float* Cp = C.p + h * C.s_h + n * C.s_n;
    // 16-byte loads need the batch's base and every row start on a 16-byte boundary (uniform over the block)
    const bool vecA = ((reinterpret_cast<uintptr_t>(Ap) & 15) == 0) && (A.ld & 3) == 0;
    const bool vecB = ((reinterpret_cast<uintptr_t>(Bp) & 15) == 0) && (B.ld & 3) == 0;
    const int m0 = blockIdx.y * BT, n0 = blockIdx.x * BT;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, q = lane & 3;
    const int wm = (warp >> 1) * 32, wn = (warp & 1) * 32;
    float acc[2][4][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = acc[i][j][2] = acc[i][j][3] = 0.f;
    const int ktiles = (K + BK - 1) / BK;
    const int kper = (ktiles + ksplit - 1) / ksplit;
    const int kt0 = split * kper, kt1 = min(ktiles, kt0 + kper);
    for (int kt = kt0; kt < kt1; ++kt) {
        const int k0 = kt * BK;
        __syncthreads();
        // operands into shared memory, four elements per thread and step: one 16-byte load where the operand's pitch and the tile
        // allow it (vecA / vecB), element loads otherwise; without the split the values are rounded to TF32 here, once, instead of
        // at every fragment load (the kernel issued 13 instructions per MMA)
#pragma unroll
        for (int i = 0; i < (BT * BK) / (128 * 4); ++i) {
            const int idx = threadIdx.x + 128 * i;
            float va[4], vb[4];
            if (!TA) {
                const int r = idx >> 3, c = (idx & 7) * 4;
                const float* src = Ap + (long)(m0 + r) * A.ld + k0 + c;
                if (vecA && m0 + r < M && k0 + c + 3 < K) {
                    const float4 t = *reinterpret_cast<const float4*>(src);
                    va[0] = t.x; va[1] = t.y; va[2] = t.z; va[3] = t.w;
                } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e) va[e] = (m0 + r < M && k0 + c + e < K) ? src[e] : 0.f;
                }
                if (!SPLIT) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) va[e] = __uint_as_float(to_tf32(va[e]));
                }
                *reinterpret_cast<float4*>(&As[r][c]) = make_float4(va[0], va[1], va[2], va[3]);
            } else {
                const int kk = idx >> 4, r = (idx & 15) * 4;
                const float* src = Ap + (long)(k0 + kk) * A.ld + m0 + r;
                if (vecA && k0 + kk < K && m0 + r + 3 < M) {
                    const float4 t = *reinterpret_cast<const float4*>(src);
                    va[0] = t.x; va[1] = t.y; va[2] = t.z; va[3] = t.w;
                } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e) va[e] = (k0 + kk < K && m0 + r + e < M) ? src[e] : 0.f;
                }
#pragma unroll
                for (int e = 0; e < 4; ++e) As[r + e][kk] = SPLIT ? va[e] : __uint_as_float(to_tf32(va[e]));
            }
            if (!TB) {
                const int r = idx >> 3, c = (idx & 7) * 4;
                const float* src = Bp + (long)(n0 + r) * B.ld + k0 + c;
                if (vecB && n0 + r < N && k0 + c + 3 < K) {
                    const float4 t = *reinterpret_cast<const float4*>(src);
                    vb[0] = t.x; vb[1] = t.y; vb[2] = t.z; vb[3] = t.w;
                } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e) vb[e] = (n0 + r < N && k0 + c + e < K) ? src[e] : 0.f;
                }
                if (!SPLIT) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) vb[e] = __uint_as_float(to_tf32(vb[e]));
                }
                *reinterpret_cast<float4*>(&Bs[r][c]) = make_float4(vb[0], vb[1], vb[2], vb[3]);
            } else {
                const int kk = idx >> 4, r = (idx & 15) * 4;
                const float* src = Bp + (long)(k0 + kk) * B.ld + n0 + r;
                if (vecB && k0 + kk < K && n0 + r + 3 < N) {
                    const float4 t = *reinterpret_cast<const float4*>(src);
                    vb[0] = t.x; vb[1] = t.y; vb[2] = t.z; vb[3] = t.w;
                } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e) vb[e] = (k0 + kk < K && n0 + r + e < N) ? src[e] : 0.f;
                }
#pragma unroll
                for (int e = 0; e < 4; ++e) Bs[r + e][kk] = SPLIT ? vb[e] : __uint_as_float(to_tf32(vb[e]));
            }
        }
        __syncthreads();
#pragma unroll
        for (int ks = 0; ks < BK / 8; ++ks) {
            uint32_t ah[2][4], al[2][4], bh[4][2], bl[4][2];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const float v[4] = {As[wm + 16 * i + g][8 * ks + q], As[wm + 16 * i + g + 8][8 * ks + q],
                                    As[wm + 16 * i + g][8 * ks + q + 4], As[wm + 16 * i + g + 8][8 * ks + q + 4]};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    ah[i][e] = SPLIT ? to_tf32(v[e]) : __float_as_uint(v[e]);
                    if (SPLIT) al[i][e] = to_tf32(v[e] - __uint_as_float(ah[i][e]));
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float v[2] = {Bs[wn + 8 * j + g][8 * ks + q], Bs[wn + 8 * j + g][8 * ks + q + 4]};
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    bh[j][e] = SPLIT ? to_tf32(v[e]) : __float_as_uint(v[e]);
                    if (SPLIT) bl[j][e] = to_tf32(v[e] - __uint_as_float(bh[j][e]));
                }
            }
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (SPLIT) {
                        mma_tf32(acc[i][j], al[i], bh[j][0], bh[j][1]);
                        mma_tf32(acc[i][j], ah[i], bl[j][0], bl[j][1]);
                    }
                    mma_tf32(acc[i][j], ah[i], bh[j][0], bh[j][1]);
                }
        }
    }
    if (kt0 >= kt1) return;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int r = m0 + wm + 16 * i + g + ((e >> 1) ? 8 : 0);
                const int c = n0 + wn + 8 * j + 2 * q + (e & 1);
                if (r < M && c < N) {
                    float* dst = Cp + (long)r * C.ld + c;
                    const float v = alpha * acc[i][j][e];
                    if (acc_mode == 0) *dst = v;
                    else if (acc_mode == 1) *dst += v;
                    else atomicAdd(dst, v);
                }
            }
}

int bgemm(bool ta, bool tb, bool split3, BMat A, BMat B, BOut C, int M, int N, int K, int Hb, int Nb, float alpha, int acc_mode,
          int ksplit, cudaStream_t st) {
    CSS_REQUIRE(M > 0 && N > 0 && K > 0 && Hb > 0 && Nb > 0 && ksplit >= 1, "bgemm: bad shape");
    CSS_REQUIRE(ksplit == 1 || acc_mode == 2, "bgemm: split-K needs the atomic epilogue");
    dim3 grid((N + BT - 1) / BT, (M + BT - 1) / BT, (unsigned)(Hb * Nb * ksplit));
    CSS_REQUIRE(grid.z <= 65535, "bgemm: too many batches (%u)", grid.z);
#define CSS_BG(TA_, TB_)                                                                                                   \
    do {                                                                                                                   \
        if (split3) bgemm_kernel<TA_, TB_, true><<<grid, 128, 0, st>>>(A, B, C, M, N, K, Nb, alpha, acc_mode, ksplit);     \
        else bgemm_kernel<TA_, TB_, false><<<grid, 128, 0, st>>>(A, B, C, M, N, K, Nb, alpha, acc_mode, ksplit);           \
    } while (0)
    if (!ta && !tb) CSS_BG(false, false);
    else if (!ta && tb) CSS_BG(false, true);
    else if (ta && !tb) CSS_BG(true, false);
    else CSS_BG(true, true);
#undef CSS_BG
    return css_check_launch("bgemm");
}

// P[b][i][:] = softmax_j((S[b][i][j] + R[b][i][i-j+T-1]) * scale), zero in the padding columns.  One warp per row.
__global__ void __launch_bounds__(256)
softmax_rel_kernel(const float* __restrict__ S, const float* __restrict__ R, float* __restrict__ P, long rows, int T, int TP, int RP) {
    const long row = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const int i = (int)(row % T);
    const float* s = S + row * TP;
    const float* r = R ? R + row * RP + i + T - 1 : nullptr;
    float v[8];
    float mx = -INFINITY;
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) {
        const int j = lane + 32 * jj;
        v[jj] = -INFINITY;
        if (j < T) {
            v[jj] = (s[j] + (r ? r[-j] : 0.f)) * 0.125f;
            mx = fmaxf(mx, v[jj]);
        }
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) {
        v[jj] = (lane + 32 * jj < T) ? expf(v[jj] - mx) : 0.f;
        sum += v[jj];
    }
    const float inv = 1.f / warp_sum(sum);
    float* p = P + row * TP;
#pragma unroll
    for (int jj = 0; jj < 8; ++jj)
        if (lane + 32 * jj < TP) p[lane + 32 * jj] = v[jj] * inv;
}

// dS = P (dP - sum_j P dP) * scale, written over dP; dR[b][i][rho] = dS[i][i + T-1 - rho] (zero elsewhere)
__global__ void __launch_bounds__(256)
ds_rel_kernel(const float* __restrict__ P, float* __restrict__ dP, float* __restrict__ dR, long rows, int T, int TP, int RP) {
    __shared__ float buf[8][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long row = (long)blockIdx.x * 8 + warp;
    if (row >= rows) return;
    const int i = (int)(row % T);
    const float* p = P + row * TP;
    float* d = dP + row * TP;
    float pv[8], dv[8];
    float dot = 0.f;
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) {
        const int j = lane + 32 * jj;
        pv[jj] = j < T ? p[j] : 0.f;
        dv[jj] = j < T ? d[j] : 0.f;
        dot += pv[jj] * dv[jj];
    }
    dot = warp_sum(dot);
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) {
        const int j = lane + 32 * jj;
        const float ds = pv[jj] * (dv[jj] - dot) * 0.125f;
        if (j < TP) d[j] = ds;
        buf[warp][j & 255] = ds;
    }
    __syncwarp();
    if (dR) {
        float* r = dR + row * RP;
        for (int rho = lane; rho < RP; rho += 32) {
            const int j = i + T - 1 - rho;
            r[rho] = (j >= 0 && j < T) ? buf[warp][j] : 0.f;
        }
    }
}

}  // namespace

size_t attention_train_scratch_floats(long N, int T, int H) {
    const int TP = (T + 31) & ~31;
    return (size_t)N * H * T * (TP + 2 * TP);      // S / dP (TP per row) + R / dR (2 TP per row)
}
size_t attention_train_probs_floats(long N, int T, int H) {
    const int TP = (T + 31) & ~31;
    return (size_t)N * H * T * TP;
}

// qkv (N*T, 3*H*64) fp32 -> out (N*T, H*64); probs (saved for the backward) and scratch sized by the helpers above
int attention_train_fwd_launch(const float* qkv, const float* pe, int pe_rows, float* out, float* probs, float* scratch, long N,
                               int T, int H, int dk, bool split3, const CssDrop* drop, cudaStream_t st) {
    CSS_REQUIRE(dk == 64 && T >= 1 && T <= 256 && N > 0 && N * H <= 65535, "attention_train: unsupported shape (T=%d dk=%d N=%ld)", T, dk, N);
    CSS_REQUIRE(!pe || T - 1 <= pe_rows / 2 - 1, "attention_train: relative positions exceed the embedding table");
    const int TP = (T + 31) & ~31, RP = 2 * TP, ld = 3 * H * dk, dm = H * dk;
    const long per = (long)T * TP, perR = (long)T * RP;
    float* S = scratch;
    float* R = scratch + (size_t)N * H * per;
    const long Nl = N;
    BMat Q = {qkv, (long)dk, (long)T * ld, ld}, Kk = {qkv + H * dk, (long)dk, (long)T * ld, ld}, V = {qkv + 2 * H * dk, (long)dk, (long)T * ld, ld};
    int rc;
    css_prof_begin(1, 6.0 * (double)N * H * T * T * dk, st);
    if (pe) {
        BMat PE = {pe + (long)(pe_rows / 2 - (T - 1)) * dk, 0, 0, dk};
        BOut Ro = {R, Nl * perR, perR, RP};
        if ((rc = bgemm(false, false, split3, Q, PE, Ro, T, 2 * T - 1, dk, H, (int)N, 1.f, 0, 1, st))) return rc;
    }
    BOut So = {S, Nl * per, per, TP};
    if ((rc = bgemm(false, false, split3, Q, Kk, So, T, T, dk, H, (int)N, 1.f, 0, 1, st))) return rc;
    const long rows = N * H * T;
    softmax_rel_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(S, pe ? R : nullptr, probs, rows, T, TP, RP);
    if ((rc = css_check_launch("softmax_rel"))) return rc;
    const float* pv = probs;
    if (drop) {
        // dropout on the probabilities (conformer.py:249): the dropped copy goes into the (now free) score buffer, `probs`
        // keeps the originals for the softmax backward.  Element index = position in the [head][n][i][TP] buffer.
        if ((rc = dropout_f32_launch(S, probs, nullptr, (long)N * H * per, *drop, 1.f, 0, st))) return rc;
        pv = S;
    }
    BMat Pm = {pv, Nl * per, per, TP};
    BOut O = {out, (long)dk, (long)T * dm, dm};
    rc = bgemm(false, true, split3, Pm, V, O, T, dk, T, H, (int)N, 1.f, 0, 1, st);
    css_prof_end(st);
    return rc;
}

// dO (N*T, H*64) -> dqkv (N*T, 3*H*64) overwritten; dpe (pe_rows, 64) accumulated
int attention_train_bwd_launch(const float* qkv, const float* pe, int pe_rows, const float* probs, const float* dO, float* dqkv,
                               float* dpe, float* scratch, long N, int T, int H, int dk, bool split3, const CssDrop* drop,
                               cudaStream_t st) {
    CSS_REQUIRE(dk == 64 && T >= 1 && T <= 256 && N > 0 && N * H <= 65535, "attention_train_bwd: unsupported shape");
    const int TP = (T + 31) & ~31, RP = 2 * TP, ld = 3 * H * dk, dm = H * dk;
    const long per = (long)T * TP, perR = (long)T * RP;
    float* dP = scratch;
    float* dR = scratch + (size_t)N * H * per;
    const long Nl = N;
    BMat Q = {qkv, (long)dk, (long)T * ld, ld}, Kk = {qkv + H * dk, (long)dk, (long)T * ld, ld}, V = {qkv + 2 * H * dk, (long)dk, (long)T * ld, ld};
    BMat dOm = {dO, (long)dk, (long)T * dm, dm};
    BMat Pm = {probs, Nl * per, per, TP}, dSm = {dP, Nl * per, per, TP};
    BOut dQ = {dqkv, (long)dk, (long)T * ld, ld}, dK = {dqkv + H * dk, (long)dk, (long)T * ld, ld}, dV = {dqkv + 2 * H * dk, (long)dk, (long)T * ld, ld};
    int rc;
    css_prof_begin(1, 14.0 * (double)N * H * T * T * dk, st);
    BOut dPo = {dP, Nl * per, per, TP};
    if (drop) {
        // dV = dropout(P)^T dO first: the dropped probabilities are re-derived into the dP buffer, which is free until dP is formed
        if ((rc = dropout_f32_launch(dP, probs, nullptr, (long)N * H * per, *drop, 1.f, 0, st))) return rc;
        BMat Pd = {dP, Nl * per, per, TP};
        if ((rc = bgemm(true, true, split3, Pd, dOm, dV, T, dk, T, H, (int)N, 1.f, 0, 1, st))) return rc;
    }
    if ((rc = bgemm(false, false, split3, dOm, V, dPo, T, T, dk, H, (int)N, 1.f, 0, 1, st))) return rc;          // dP = dO V^T
    if (drop && (rc = dropout_f32_launch(dP, dP, nullptr, (long)N * H * per, *drop, 1.f, 0, st))) return rc;     // through the mask
    const long rows = N * H * T;
    ds_rel_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(probs, dP, pe ? dR : nullptr, rows, T, TP, RP);
    if ((rc = css_check_launch("ds_rel"))) return rc;
    if (!drop && (rc = bgemm(true, true, split3, Pm, dOm, dV, T, dk, T, H, (int)N, 1.f, 0, 1, st))) return rc;    // dV = P^T dO
    if ((rc = bgemm(true, true, split3, dSm, Q, dK, T, dk, T, H, (int)N, 1.f, 0, 1, st))) return rc;              // dK = dS^T Q
    if ((rc = bgemm(false, true, split3, dSm, Kk, dQ, T, dk, T, H, (int)N, 1.f, 0, 1, st))) return rc;            // dQ = dS K
    if (pe) {
        const float* pew = pe + (long)(pe_rows / 2 - (T - 1)) * dk;
        BMat dRm = {dR, Nl * perR, perR, RP};
        BMat PEw = {pew, 0, 0, dk};
        if ((rc = bgemm(false, true, split3, dRm, PEw, dQ, T, dk, 2 * T - 1, H, (int)N, 1.f, 1, 1, st))) return rc;   // dQ += dR PE_win
        // dPE_win[rho][d] += sum over (n, i) of dR[h][n][i][rho] Q_h[(n,i)][d]: one long contraction per head, split over the SMs
        BMat dRall = {dR, Nl * perR, 0, RP};
        BMat Qall = {qkv, (long)dk, 0, ld};
        BOut dPE = {dpe + (long)(pe_rows / 2 - (T - 1)) * dk, 0, 0, dk};
        const long kk = N * T;
        int ksplit = (int)((kk + 8 * BK - 1) / (8 * BK));
        if (ksplit > 128) ksplit = 128;
        if ((rc = bgemm(true, true, split3, dRall, Qall, dPE, 2 * T - 1, dk, (int)kk, H, 1, 1.f, 2, ksplit, st))) return rc;
    }
    css_prof_end(st);
    return 0;
}
