// K9-K11: mask-based Souden MVDR per chunk (css/executor/beamformer.py:126-183 and the
// asteroid.dsp.beamforming arithmetic it calls: compute_scm + SoudenMVDRBeamformer, restated in
// SURVEY.md appendix B).  Spatial covariances are accumulated ONCE per (chunk, bin) for the two
// speech masks and the noise mask, the 7x7 complex systems are solved in fp64 with partial
// pivoting, and the beamformer is applied in the same kernel.
//
// Centred-STFT frame t of chunk b (Q1: one-frame skew w.r.t. the mask frames):
//   t == 0 -> edge[d][b][k]   (reflect-padded frame),   t >= 1 -> spec[d][b*hop_frames + t - 1][k]
// stitched: [3][L][KP] (speech0, speech1, noise), chunk b uses frames [b*hop_frames, +T).
// Y: [B][2][T][KP] complex64.
#include "common.cuh"
#include "tma_map.h"

#define MV_M 7
#define MV_TRI 28          // upper triangle incl. diagonal
#define MV_E (2 * MV_TRI)  // packed reals per covariance: (re, im) of (i, j >= i), row-major
#define MV_BINS 32

// ---- kernel 1: masked covariances.  grid (9 bin tiles, B); block (32 bins, 4 pair groups).
// The outer product x_i conj(x_j) does not depend on the mask, so one thread forms it once and feeds the three masks'
// accumulators (42 per thread); the 28 pairs (i, j >= i) are dealt to the four warps by rows {0}, {1,6}, {2,5}, {3,4}
// -- 7 pairs each.  The next frame's 7 channels and 3 mask values are requested before the current frame is consumed.
// scm[b][mask][e][KP] fp32 (normalised by the mask sum, compute_scm(normalize=True)); energy slot 1 = sum |mask x0|^2.
// row I of the upper triangle: pairs (I, j >= I) into acc[.][Q0 ...]
template <int I, int Q0>
__device__ __forceinline__ void scm_row(const float2 (&x)[MV_M], const float (&w)[3], float2 (&acc)[3][7]) {
    const float2 xi = x[I];
    const float2 xs = make_float2(xi.y, -xi.x);
#pragma unroll
    for (int j = I; j < MV_M; ++j) {
        // x_i * conj(x_j) = (xr_i xr_j + xi_i xi_j, xi_i xr_j - xr_i xi_j)
        const float2 pr = __ffma2_rn(xi, make_float2(x[j].x, x[j].x), __fmul2_rn(xs, make_float2(x[j].y, x[j].y)));
#pragma unroll
        for (int m = 0; m < 3; ++m) acc[m][Q0 + j - I] = __ffma2_rn(make_float2(w[m], w[m]), pr, acc[m][Q0 + j - I]);
    }
}

// Frames reach the four warps through a ring of SC_ST stages of SC_F frames in shared memory, filled by the bulk-copy engine
// (cp.async.bulk, one copy per row piece, completion on the stage's mbarrier; 7 channels x 32 bins complex + 3 masks x 32 bins per
// frame = 2,176 B): the first version had every warp load its own
// operands one frame ahead and spent its time waiting for one round trip per frame (15 % of the DRAM rate at 20 warps per SM);
// now SC_ST - 1 stages are in flight per CTA and the four warps share one copy of the frame.
#define SC_F 5
#define SC_ST 4
#define SC_XF (MV_M * MV_BINS)            // float2 per frame
#define SC_WF (3 * MV_BINS)               // float per frame

__device__ __forceinline__ uint32_t sc_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void sc_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(sc_smem(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void sc_mbar_expect(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sc_smem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void sc_mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.b32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(sc_smem(bar)), "r"(parity) : "memory");
    } while (!done);
}
// one row piece (a multiple of 16 bytes) global -> shared by the bulk-copy engine, completion counted on the stage's mbarrier
__device__ __forceinline__ void sc_bulk(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(sc_smem(dst)), "l"(src), "r"(bytes), "r"(sc_smem(bar)) : "memory");
}

// one box of a 3-D tensor map global -> shared (TMA), completion counted on the stage's mbarrier
__device__ __forceinline__ void sc_tma3(void* dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(sc_smem(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(sc_smem(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// Tensor maps of the covariance kernel (fp32 views; a complex bin = two floats):
//   x5 / x1: spectrum (2 KP, channels, frames), boxes of 32 bins x 7 channels x SC_F frames / x 1 frame
//   e1:      chunk-start edge frames (2 KP, channels, chunks), box 32 bins x 7 channels x 1
//   w5:      stitched masks (KP, 3, L), box 32 bins x 3 masks x SC_F frames
// A box lands in shared memory frame-major, [frame][channel or mask][bin]: the layout of the stage buffers.
struct ScmMaps { CUtensorMap x5, x1, e1, w5; };

template <int G>
__device__ __forceinline__ void scm_group(const ScmMaps& maps, int hop_frames, int T, float* __restrict__ scm,
                                          float* __restrict__ energy, float2* Xs, float* Ws, uint64_t* full) {
    constexpr int R0 = G, R1 = (G == 0) ? -1 : MV_M - G;
    const long b = blockIdx.y;
    const int lane = threadIdx.x, tid = threadIdx.y * MV_BINS + threadIdx.x;
    const int k0 = blockIdx.x * MV_BINS, k = k0 + lane;
    float2 acc[3][7];                        // (re, im) of the thread's 7 pairs, per mask: one packed FMA (FFMA2) each
#pragma unroll
    for (int m = 0; m < 3; ++m)
#pragma unroll
        for (int i = 0; i < 7; ++i) acc[m][i] = make_float2(0.f, 0.f);
    float msum[3] = {0.f, 0.f, 0.f}, e_mask[2] = {0.f, 0.f};

    // Stage st = frames [st * SC_F, +SC_F) of the chunk.  Frame t >= 1 is global frame b * hop + t - 1, frame 0 the reflect-padded
    // edge frame of the chunk, so stage 0 is one edge box + SC_F - 1 single-frame boxes and every other stage ONE spectrum box;
    // the masks are one box per stage.  (Earlier forms: 16-byte cp.async pieces from all threads -- their address arithmetic was
    // 40 % of the kernel's instructions; then one bulk copy per 256-byte row piece, 50 per stage -- the copy engine then took
    // ~20 cycles per small copy and the warps waited at the stage barrier.)  Bins past the row (last bin tile) and frames past the
    // spectrogram are out-of-bounds box elements: they read as zero and are never used.
    constexpr uint32_t kStageBytes = SC_F * (SC_XF * 8 + SC_WF * 4);
    auto load_stage = [&](int st) {
        if (lane != 0) return;
        const int buf = st % SC_ST;
        const int g0 = (int)(b * hop_frames) + st * SC_F;              // global mask frame of the stage's first frame
        sc_mbar_expect(&full[buf], kStageBytes);
        float2* xd = Xs + (buf * SC_F) * SC_XF;
        if (st == 0) {
            sc_tma3(xd, &maps.e1, 2 * k0, 0, (int)b, &full[buf]);
#pragma unroll
            for (int f = 1; f < SC_F; ++f) sc_tma3(xd + f * SC_XF, &maps.x1, 2 * k0, 0, g0 + f - 1, &full[buf]);
        } else {
            sc_tma3(xd, &maps.x5, 2 * k0, 0, g0 - 1, &full[buf]);
        }
        sc_tma3(Ws + (buf * SC_F) * SC_WF, &maps.w5, k0, 0, g0, &full[buf]);
    };

    const int n_st = (T + SC_F - 1) / SC_F;
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < SC_ST; ++i) sc_mbar_init(&full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (G == 0) {
#pragma unroll
        for (int s0 = 0; s0 < SC_ST - 1; ++s0)
            if (s0 < n_st) load_stage(s0);
    }
    for (int st = 0; st < n_st; ++st) {
        const int buf = st % SC_ST, nf = min(SC_F, T - st * SC_F);
        sc_mbar_wait(&full[buf], (uint32_t)(st / SC_ST) & 1u);
        __syncthreads();                     // stage st - 1 has been consumed by everyone: its buffer takes stage st + SC_ST - 1
        if (G == 0 && st + SC_ST - 1 < n_st) load_stage(st + SC_ST - 1);
        for (int f = 0; f < nf; ++f) {
            float2 x[MV_M];
            float w[3];
            const float2* xf = Xs + (buf * SC_F + f) * SC_XF + lane;
            const float* wf = Ws + (buf * SC_F + f) * SC_WF + lane;
#pragma unroll
            for (int d = 0; d < MV_M; ++d) x[d] = xf[d * MV_BINS];
#pragma unroll
            for (int m = 0; m < 3; ++m) w[m] = wf[m * MV_BINS];
#pragma unroll
            for (int m = 0; m < 3; ++m) msum[m] += w[m];
            if (G == 0) {
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    const float ar = w[m] * x[0].x, ai = w[m] * x[0].y;
                    e_mask[m] += ar * ar + ai * ai;
                }
            }
            scm_row<R0, 0>(x, w, acc);
            if (R1 >= 0) scm_row<(R1 >= 0 ? R1 : 0), MV_M - R0>(x, w, acc);
        }
    }
    if (k < CSS_NBINS) {
#pragma unroll
        for (int m = 0; m < 3; ++m) {
            float* dst = scm + ((b * 3 + m) * (long)MV_E) * CSS_KP + k;
            const float inv = 1.f / msum[m];
            int q = 0;
#pragma unroll
            for (int rr = 0; rr < 2; ++rr) {
                const int i = rr == 0 ? R0 : R1;
                if (i < 0) continue;
                const int e0 = 2 * (i * MV_M - (i * (i - 1)) / 2);        // packed slot of pair (i, i)
#pragma unroll
                for (int j = i; j < MV_M; ++j) {
                    dst[(long)(e0 + 2 * (j - i)) * CSS_KP] = acc[m][q].x * inv;
                    dst[(long)(e0 + 2 * (j - i) + 1) * CSS_KP] = acc[m][q].y * inv;
                    ++q;
                }
            }
        }
    }
    if (G == 0) {
#pragma unroll
        for (int m = 0; m < 2; ++m) {
            const float e = warp_sum(k < CSS_NBINS ? e_mask[m] : 0.f);
            if (lane == 0) energy[((b * 2 + m) * gridDim.x + blockIdx.x) * 2 + 1] = e;
        }
    }
}

__global__ void __launch_bounds__(128)
mvdr_scm_kernel(const __grid_constant__ ScmMaps maps, int hop_frames, int T, float* __restrict__ scm, float* __restrict__ energy) {
    __shared__ __align__(128) float2 Xs[SC_ST * SC_F * SC_XF];
    __shared__ __align__(128) float Ws[SC_ST * SC_F * SC_WF];
    __shared__ __align__(8) uint64_t full[SC_ST];
    switch (threadIdx.y) {      // warp-uniform; every arm runs the same copy / barrier sequence
        case 0: scm_group<0>(maps, hop_frames, T, scm, energy, Xs, Ws, full); break;
        case 1: scm_group<1>(maps, hop_frames, T, scm, energy, Xs, Ws, full); break;
        case 2: scm_group<2>(maps, hop_frames, T, scm, energy, Xs, Ws, full); break;
        default: scm_group<3>(maps, hop_frames, T, scm, energy, Xs, Ws, full); break;
    }
}

// ---- kernel 2: N = solve(Phi_n, Phi_s) in complex128, w = (N / (tr N + eps))[:, 0].
// Eight lanes share one 7x7 system (lane r owns row r of [Phi_n | Phi_s]); Gauss-Jordan with partial pivoting
// (largest |re|+|im| among the rows not yet used as pivot, as LAPACK's izamax) entirely in registers: the pivot
// row is broadcast with shuffles, no row is ever moved.  grid (9, B, 2 speakers); block 256 = 32 systems.
struct cd { double x, y; };
__device__ __forceinline__ cd cd_mul(cd a, cd b) { return {a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x}; }
__device__ __forceinline__ cd cd_sub(cd a, cd b) { return {a.x - b.x, a.y - b.y}; }
__device__ __forceinline__ cd cd_div(cd a, cd b) {
    // Smith's algorithm
    if (fabs(b.x) >= fabs(b.y)) {
        const double r = b.y / b.x, den = b.x + b.y * r;
        return {(a.x + a.y * r) / den, (a.y - a.x * r) / den};
    }
    const double r = b.x / b.y, den = b.x * r + b.y;
    return {(a.x * r + a.y) / den, (a.y * r - a.x) / den};
}
__device__ __forceinline__ cd cd_shfl(cd v, int src) {
    return {__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src)};
}
// element (r, j) of a Hermitian matrix packed as the upper triangle in shared memory [e][32 bins]
__device__ __forceinline__ cd herm_at(const float* __restrict__ m, int r, int j, int bin) {
    const int i0 = r < j ? r : j, j0 = r < j ? j : r;
    const int q = i0 * MV_M - (i0 * (i0 - 1)) / 2 + (j0 - i0);
    const float re = m[(2 * q) * MV_BINS + bin], im = m[(2 * q + 1) * MV_BINS + bin];
    return {(double)re, r <= j ? (double)im : -(double)im};
}

__global__ void __launch_bounds__(256)
mvdr_solve_kernel(const float* __restrict__ scm, float2* __restrict__ wout /* [B][2][7][KP] */) {
    __shared__ float sn[MV_E * MV_BINS], ss[MV_E * MV_BINS];
    const long b = blockIdx.y;
    const int spk = blockIdx.z;
    const int k0 = blockIdx.x * MV_BINS;
    for (int i = threadIdx.x; i < MV_E * MV_BINS; i += blockDim.x) {
        const int e = i >> 5, kk = i & 31;
        const bool ok = k0 + kk < CSS_NBINS;
        sn[i] = ok ? scm[((b * 3 + 2) * (long)MV_E + e) * CSS_KP + k0 + kk] : 0.f;
        ss[i] = ok ? scm[((b * 3 + spk) * (long)MV_E + e) * CSS_KP + k0 + kk] : 0.f;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, r = lane & 7;
    const int bin = (threadIdx.x >> 5) * 4 + (lane >> 3);        // 0..31
    const int grp = lane & ~7;
    const bool real_row = r < MV_M;
    cd A[MV_M], Bm[MV_M];
#pragma unroll
    for (int j = 0; j < MV_M; ++j) {
        A[j] = real_row ? herm_at(sn, r, j, bin) : cd{0.0, 0.0};
        Bm[j] = real_row ? herm_at(ss, r, j, bin) : cd{0.0, 0.0};
    }
    bool used = !real_row;      // the dummy eighth lane never pivots
    int my_col = -1;
#pragma unroll
    for (int c = 0; c < MV_M; ++c) {
        double v = used ? -1.0 : fabs(A[c].x) + fabs(A[c].y);
        int p = lane;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            const double vo = __shfl_xor_sync(0xffffffffu, v, o);
            const int po = __shfl_xor_sync(0xffffffffu, p, o);
            if (vo > v || (vo == v && po < p)) { v = vo; p = po; }
        }
        const bool is_piv = lane == p;
        const cd pivc = cd_shfl(A[c], p);
        const cd f = (is_piv || !real_row) ? cd{0.0, 0.0} : cd_div(A[c], pivc);
#pragma unroll
        for (int j = 0; j < MV_M; ++j) {
            if (j > c) { const cd pj = cd_shfl(A[j], p); A[j] = cd_sub(A[j], cd_mul(f, pj)); }
            const cd pb = cd_shfl(Bm[j], p);
            Bm[j] = cd_sub(Bm[j], cd_mul(f, pb));
        }
        if (!is_piv) A[c] = cd{0.0, 0.0};
        if (is_piv) { used = true; my_col = c; }
    }
    // the lane that pivoted column c now holds row c of N: N[c][j] = Bm[j] / A[c]
    cd diag = {1.0, 0.0}, n0 = {0.0, 0.0}, ncc = {0.0, 0.0};
#pragma unroll
    for (int c = 0; c < MV_M; ++c)
        if (my_col == c) { diag = A[c]; ncc = Bm[c]; }
    if (real_row) { n0 = cd_div(Bm[0], diag); ncc = cd_div(ncc, diag); }
    // back to complex64 (asteroid casts the solve result to the input dtype), trace in fp32, divide
    float tr_r = real_row ? (float)ncc.x : 0.f, tr_i = real_row ? (float)ncc.y : 0.f;
    // sum in column order 0..6 like the sequential fp32 trace: gather each column's term from its owner lane
    float sr = 0.f, si = 0.f;
#pragma unroll
    for (int c = 0; c < MV_M; ++c) {
        const unsigned m = __ballot_sync(0xffffffffu, my_col == c);
        const int owner = __ffs((m >> grp) & 0xffu) - 1 + grp;
        sr += __shfl_sync(0xffffffffu, tr_r, owner);
        si += __shfl_sync(0xffffffffu, tr_i, owner);
    }
    sr += 1e-8f;
    const float den = sr * sr + si * si;
    const int k = k0 + bin;
    if (real_row && k < CSS_NBINS) {
        const float nr = (float)n0.x, ni = (float)n0.y;
        wout[((b * 2 + spk) * MV_M + my_col) * (long)CSS_KP + k] = make_float2((nr * sr + ni * si) / den, (ni * sr - nr * si) / den);
    }
}

// ---- kernel 3: y[t] = sum_m conj(w_m) x_m[t];  grid (9, B); block (32 bins, 2 speakers); energy slot 0 = sum |y|^2
// The frames arrive through the same TMA ring as in the covariance kernel (one box of 32 bins x 7 channels x 5 frames per stage,
// shared by the two speakers' warps): with every thread loading its own operands the kernel waited for one round trip per
// group of frames (ncu: long_scoreboard 23 of 27 warp-cycles per issue).
__global__ void __launch_bounds__(64)
mvdr_apply_kernel(const __grid_constant__ ScmMaps maps, int hop_frames, int T, const float2* __restrict__ wv, float2* __restrict__ Y,
                  float* __restrict__ energy) {
    __shared__ __align__(128) float2 Xs[SC_ST * SC_F * SC_XF];
    __shared__ __align__(8) uint64_t full[SC_ST];
    const long b = blockIdx.y;
    const int lane = threadIdx.x, mk = threadIdx.y, tid = mk * MV_BINS + lane;
    const int k0 = blockIdx.x * MV_BINS, k = k0 + lane;
    const bool act = k < CSS_NBINS;
    float e_y = 0.f;
    float2* yrow = Y + ((b * 2 + mk) * (long)T) * CSS_KP + k;
    // y = sum_d conj(w_d) x_d = (P.x + Q.y, P.y - Q.x) with P = sum_d wr_d * x_d, Q = sum_d wi_d * x_d: two packed fp32
    // FMAs (FFMA2) per channel
    float2 wx[MV_M], wy[MV_M];
#pragma unroll
    for (int i = 0; i < MV_M; ++i) {
        const float2 w = act ? wv[((b * 2 + mk) * MV_M + i) * (long)CSS_KP + k] : make_float2(0.f, 0.f);
        wx[i] = make_float2(w.x, w.x);
        wy[i] = make_float2(w.y, w.y);
    }
    constexpr uint32_t kStageBytes = SC_F * SC_XF * 8;
    auto load_stage = [&](int st) {
        const int buf = st % SC_ST;
        const int g0 = (int)(b * hop_frames) + st * SC_F;
        sc_mbar_expect(&full[buf], kStageBytes);
        float2* xd = Xs + (buf * SC_F) * SC_XF;
        if (st == 0) {
            sc_tma3(xd, &maps.e1, 2 * k0, 0, (int)b, &full[buf]);
#pragma unroll
            for (int f = 1; f < SC_F; ++f) sc_tma3(xd + f * SC_XF, &maps.x1, 2 * k0, 0, g0 + f - 1, &full[buf]);
        } else {
            sc_tma3(xd, &maps.x5, 2 * k0, 0, g0 - 1, &full[buf]);
        }
    };
    const int n_st = (T + SC_F - 1) / SC_F;
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < SC_ST; ++i) sc_mbar_init(&full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
#pragma unroll
        for (int s0 = 0; s0 < SC_ST - 1; ++s0)
            if (s0 < n_st) load_stage(s0);
    }
    for (int st = 0; st < n_st; ++st) {
        const int buf = st % SC_ST, nf = min(SC_F, T - st * SC_F);
        sc_mbar_wait(&full[buf], (uint32_t)(st / SC_ST) & 1u);
        __syncthreads();                     // stage st - 1 has been consumed by both warps
        if (tid == 0 && st + SC_ST - 1 < n_st) load_stage(st + SC_ST - 1);
        for (int f = 0; f < nf; ++f) {
            const int t = st * SC_F + f;
            const float2* xf = Xs + (buf * SC_F + f) * SC_XF + lane;
            float2 P = make_float2(0.f, 0.f), Q = make_float2(0.f, 0.f);
#pragma unroll
            for (int d = 0; d < MV_M; ++d) {
                const float2 x = xf[d * MV_BINS];
                P = __ffma2_rn(wx[d], x, P);
                Q = __ffma2_rn(wy[d], x, Q);
            }
            const float yr = P.x + Q.y, yi = P.y - Q.x;      // zero for the padding bins (zero weights)
            if (k < CSS_KP) yrow[(long)t * CSS_KP] = make_float2(yr, yi);
            if (act) e_y += yr * yr + yi * yi;
        }
    }
    e_y = warp_sum(e_y);
    if (lane == 0) energy[((b * 2 + mk) * gridDim.x + blockIdx.x) * 2] = e_y;
}

extern "C" int css_mvdr(const void* spec, const void* edge, long frames, long n_chunks, const float* stitched, long L,
                        int hop_frames, int T, void* Y, float* energy, cudaStream_t st) {
    CSS_REQUIRE(n_chunks > 0 && T > 1, "css_mvdr: empty input");
    CSS_REQUIRE(L >= (n_chunks - 1) * hop_frames + T, "css_mvdr: stitched masks too short (L=%ld)", L);
    CSS_REQUIRE(frames >= (n_chunks - 1) * hop_frames + T - 1, "css_mvdr: spectrogram too short (frames=%ld)", frames);
    CSS_REQUIRE(n_chunks <= 65535, "css_mvdr: more than 65535 chunks per call");
    // stream-ordered scratch: covariances [B][3][56][KP] fp32 and beamformer weights [B][2][7][KP] complex64
    const size_t scm_bytes = (size_t)n_chunks * 3 * MV_E * CSS_KP * 4, w_bytes = (size_t)n_chunks * 2 * MV_M * CSS_KP * 8;
    void* scratch = nullptr;
    cudaError_t e = cudaMallocAsync(&scratch, scm_bytes + w_bytes, st);
    CSS_REQUIRE(e == cudaSuccess, "css_mvdr: scratch allocation of %zu B failed: %s", scm_bytes + w_bytes, cudaGetErrorString(e));
    float* scm = static_cast<float*>(scratch);
    float2* wv = reinterpret_cast<float2*>(static_cast<char*>(scratch) + scm_bytes);
    const unsigned nblk = (CSS_NBINS + MV_BINS - 1) / MV_BINS;
    CssProfScope prof(9, (double)n_chunks * T * CSS_NBINS * (7 * 8.0 + 3 * 4.0 + 2 * 8.0), st);
    ScmMaps maps;
    {
        int rc0;
        const unsigned long long KP2 = 2ULL * CSS_KP;
        const unsigned long long dx[3] = {KP2, MV_M, (unsigned long long)frames}, sx[2] = {(unsigned long long)frames * CSS_KP * 8ULL, CSS_KP * 8ULL};
        const unsigned long long de[3] = {KP2, MV_M, (unsigned long long)n_chunks}, se[2] = {(unsigned long long)n_chunks * CSS_KP * 8ULL, CSS_KP * 8ULL};
        const unsigned long long dw[3] = {CSS_KP, 3, (unsigned long long)L}, sw[2] = {(unsigned long long)L * CSS_KP * 4ULL, CSS_KP * 4ULL};
        const unsigned b5[3] = {2 * MV_BINS, MV_M, SC_F}, b1[3] = {2 * MV_BINS, MV_M, 1}, bw[3] = {MV_BINS, 3, SC_F};
        if ((rc0 = css_make_f32_map3(&maps.x5, spec, dx, sx, b5)) || (rc0 = css_make_f32_map3(&maps.x1, spec, dx, sx, b1)) ||
            (rc0 = css_make_f32_map3(&maps.e1, edge, de, se, b1)) || (rc0 = css_make_f32_map3(&maps.w5, stitched, dw, sw, bw))) {
            cudaFreeAsync(scratch, st);
            return rc0;
        }
    }
    mvdr_scm_kernel<<<dim3(nblk, (unsigned)n_chunks), dim3(MV_BINS, 4), 0, st>>>(maps, hop_frames, T, scm, energy);
    int rc = css_check_launch("css_mvdr(scm)");
    if (!rc) {
        mvdr_solve_kernel<<<dim3(nblk, (unsigned)n_chunks, 2), 256, 0, st>>>(scm, wv);
        rc = css_check_launch("css_mvdr(solve)");
    }
    if (!rc) {
        mvdr_apply_kernel<<<dim3(nblk, (unsigned)n_chunks), dim3(MV_BINS, 2), 0, st>>>(maps, hop_frames, T, wv, (float2*)Y, energy);
        rc = css_check_launch("css_mvdr(apply)");
    }
    cudaFreeAsync(scratch, st);
    return rc;
}
extern "C" int css_mvdr_energy_blocks(void) { return (CSS_NBINS + MV_BINS - 1) / MV_BINS; }

// ---------------------------------------------------------------------------------------------
// K11: per-chunk RMS rescale (beamformer.py:165-183) and dedup gain (beamformer.py:68-80).
// scal[b] = {mvdr_rms0, mask_rms0, mvdr_rms1, mask_rms1, mode}; mode 1: attenuate spk1, 2: spk0.
// ---------------------------------------------------------------------------------------------
__global__ void chunk_scalars_kernel(const float* __restrict__ energy, long n_chunks, int nblk, int T, int rescale,
                                     float* __restrict__ scal) {
    const long b = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_chunks) return;
    float pw[2];
    for (int s = 0; s < 2; ++s) {
        float ey = 0.f, em = 0.f;
        for (int i = 0; i < nblk; ++i) {
            ey += energy[((b * 2 + s) * nblk + i) * 2];
            em += energy[((b * 2 + s) * nblk + i) * 2 + 1];
        }
        const float n = (float)CSS_NBINS * T;
        const float r_y = sqrtf(ey / n), r_m = sqrtf(em / n);
        scal[b * 5 + 2 * s] = r_y;
        scal[b * 5 + 2 * s + 1] = r_m;
        const float g = rescale ? (r_m / r_y) : 1.f;
        pw[s] = 10.f * log10f(ey * g * g);
    }
    float mode = 0.f;
    if (pw[0] - pw[1] > 15.f) mode = 1.f;
    else if (pw[1] - pw[0] > 15.f) mode = 2.f;
    scal[b * 5 + 4] = mode;
}

__global__ void __launch_bounds__(256)
mvdr_finalize_kernel(float2* __restrict__ Y, const float* __restrict__ scal, int T, int rescale, int dedup) {
    const long b = blockIdx.y;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= T * CSS_KP) return;
    const int k = e % CSS_KP;
    if (k >= CSS_NBINS) return;
    float2* y0 = Y + (b * 2) * (long)T * CSS_KP + e;
    float2* y1 = y0 + (long)T * CSS_KP;
    float2 a = *y0, c = *y1;
    const float* sc = scal + b * 5;
    if (rescale) {
        a.x = a.x / sc[0] * sc[1]; a.y = a.y / sc[0] * sc[1];
        c.x = c.x / sc[2] * sc[3]; c.y = c.y / sc[2] * sc[3];
    }
    const int mode = dedup ? (int)sc[4] : 0;
    if (mode) {
        const float m0 = sqrtf(a.x * a.x + a.y * a.y), m1 = sqrtf(c.x * c.x + c.y * c.y);
        const float mx = fmaxf(m0, m1);
        if (mode == 1) { const float g = fmaxf(m1 / mx, 0.01f); c.x *= g; c.y *= g; }
        else           { const float g = fmaxf(m0 / mx, 0.01f); a.x *= g; a.y *= g; }
    }
    *y0 = a;
    *y1 = c;
}

extern "C" int css_mvdr_finalize(void* Y, const float* energy, long n_chunks, int T, int rescale, int dedup,
                                 float* scal, cudaStream_t st) {
    CSS_REQUIRE(n_chunks > 0 && n_chunks <= 65535, "css_mvdr_finalize: chunk count out of range");
    const int nblk = (CSS_NBINS + MV_BINS - 1) / MV_BINS;
    CssProfScope prof(9, 0.0, st);   // second pass over Y: time counted with the MVDR, bytes already counted there
    chunk_scalars_kernel<<<(unsigned)((n_chunks + 127) / 128), 128, 0, st>>>(energy, n_chunks, nblk, T, rescale, scal);
    int rc = css_check_launch("chunk_scalars");
    if (rc) return rc;
    if (!rescale && !dedup) return 0;
    dim3 grid((T * CSS_KP + 255) / 256, (unsigned)n_chunks);
    mvdr_finalize_kernel<<<grid, 256, 0, st>>>((float2*)Y, scal, T, rescale, dedup);
    return css_check_launch("css_mvdr_finalize");
}

// ---------------------------------------------------------------------------------------------
// Beamformer.normalize_scale as a stand-alone call (css/executor/beamformer.py:165-183):
//   out[b] = result[b] * rms(mask[b] * ref[b]) / rms(result[b])   over the n = F*T elements of chunk b.
// One CTA per chunk: both energies in one pass (fp32 terms, fp64 accumulation), then the scaled copy.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512)
normalize_scale_kernel(const float2* __restrict__ result, const float2* __restrict__ ref, const float* __restrict__ mask, long n,
                       float2* __restrict__ out) {
    __shared__ double red[2][16];
    __shared__ float gain_s;
    const long b = blockIdx.x;
    const float2* r = result + b * n;
    const float2* x = ref + b * n;
    const float* m = mask + b * n;
    double ey = 0.0, em = 0.0;
    for (long i = threadIdx.x; i < n; i += blockDim.x) {
        const float2 y = r[i], z = x[i];
        const float w = m[i];
        ey += (double)(y.x * y.x + y.y * y.y);
        em += (double)((w * z.x) * (w * z.x) + (w * z.y) * (w * z.y));
    }
    ey = warp_sum_d(ey);
    em = warp_sum_d(em);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = ey; red[1][threadIdx.x >> 5] = em; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double sy = 0.0, sm = 0.0;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { sy += red[0][i]; sm += red[1][i]; }
        gain_s = sqrtf((float)(sm / (double)n)) / sqrtf((float)(sy / (double)n));
    }
    __syncthreads();
    const float g = gain_s;
    for (long i = threadIdx.x; i < n; i += blockDim.x) {
        const float2 y = r[i];
        out[b * n + i] = make_float2(y.x * g, y.y * g);
    }
}

extern "C" int css_normalize_scale(const void* result, const void* ref, const float* mask, long n_chunks, long n, void* out,
                                   cudaStream_t st) {
    CSS_REQUIRE(n_chunks > 0 && n > 0, "css_normalize_scale: empty input");
    normalize_scale_kernel<<<(unsigned)n_chunks, 512, 0, st>>>((const float2*)result, (const float2*)ref, mask, n, (float2*)out);
    return css_check_launch("css_normalize_scale");
}
