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

#define MV_M 7
#define MV_TRI 28          // upper triangle incl. diagonal
#define MV_BINS 32

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

__device__ __forceinline__ float2 load_x(const float2* __restrict__ spec, const float2* __restrict__ edge, long frames,
                                         long n_chunks, long b, int hop_frames, int d, int t, int k) {
    return (t == 0) ? edge[((long)d * n_chunks + b) * CSS_KP + k]
                    : spec[((long)d * frames + b * hop_frames + t - 1) * CSS_KP + k];
}

// grid (ceil(257/32), B); block (32 bins, 3 masks)
__global__ void __launch_bounds__(96)
mvdr_kernel(const float2* __restrict__ spec, const float2* __restrict__ edge, long frames, long n_chunks,
            const float* __restrict__ stitched, long L, int hop_frames, int T, float2* __restrict__ Y,
            float* __restrict__ energy /* [B][2][nblk][2] : sum|y|^2, sum|mask*x0|^2 */) {
    __shared__ float scm[3][MV_BINS][2 * MV_TRI];     // per mask, per bin: 7 real diag + 21 complex
    __shared__ float2 wv[2][MV_BINS][MV_M];
    __shared__ float esum[2][2];
    const long b = blockIdx.y;
    const int lane = threadIdx.x, mk = threadIdx.y;
    const int k = blockIdx.x * MV_BINS + lane;
    const bool act = k < CSS_NBINS;
    if (lane < 2 && mk < 2) esum[mk][lane] = 0.f;
    // ---- phase 1: masked covariance, 49 real accumulators per thread ----
    float acc[2 * MV_TRI];
#pragma unroll
    for (int i = 0; i < 2 * MV_TRI; ++i) acc[i] = 0.f;
    float msum = 0.f, e_mask = 0.f, e_y = 0.f;
    if (act) {
        const float* mrow = stitched + ((long)mk * L + b * hop_frames) * CSS_KP + k;
        for (int t = 0; t < T; ++t) {
            const float w = mrow[(long)t * CSS_KP];
            float2 x[MV_M];
#pragma unroll
            for (int d = 0; d < MV_M; ++d) x[d] = load_x(spec, edge, frames, n_chunks, b, hop_frames, d, t, k);
            msum += w;
            if (mk < 2) {
                const float ar = w * x[0].x, ai = w * x[0].y;
                e_mask += ar * ar + ai * ai;
            }
            int q = 0;
#pragma unroll
            for (int i = 0; i < MV_M; ++i) {
                const float wr = w * x[i].x, wi = w * x[i].y;
#pragma unroll
                for (int j = i; j < MV_M; ++j) {
                    // (w x_i) * conj(x_j)
                    acc[q++] += wr * x[j].x + wi * x[j].y;
                    acc[q++] += wi * x[j].x - wr * x[j].y;
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 2 * MV_TRI; ++i) scm[mk][lane][i] = acc[i] / msum;   // compute_scm(normalize=True)
    }
    __syncthreads();
    // ---- phase 2: N = solve(Phi_n, Phi_s) in complex128; w = (N / (tr N + eps))[:, 0] ----
    if (act && mk < 2) {
        cd A[MV_M][MV_M], Bm[MV_M][MV_M];
        int q = 0;
        for (int i = 0; i < MV_M; ++i)
            for (int j = i; j < MV_M; ++j) {
                const float nr = scm[2][lane][q], ni = scm[2][lane][q + 1];
                const float sr = scm[mk][lane][q], si = scm[mk][lane][q + 1];
                q += 2;
                A[i][j] = {(double)nr, (double)ni};
                Bm[i][j] = {(double)sr, (double)si};
                if (j != i) {
                    A[j][i] = {(double)nr, -(double)ni};
                    Bm[j][i] = {(double)sr, -(double)si};
                }
            }
        // LU with partial pivoting (pivot by |re|+|im| as LAPACK izamax), applied to all 7 RHS
        for (int c = 0; c < MV_M; ++c) {
            int p = c;
            double best = fabs(A[c][c].x) + fabs(A[c][c].y);
            for (int r = c + 1; r < MV_M; ++r) {
                const double v = fabs(A[r][c].x) + fabs(A[r][c].y);
                if (v > best) { best = v; p = r; }
            }
            if (p != c)
                for (int j = 0; j < MV_M; ++j) {
                    cd t = A[c][j]; A[c][j] = A[p][j]; A[p][j] = t;
                    t = Bm[c][j]; Bm[c][j] = Bm[p][j]; Bm[p][j] = t;
                }
            for (int r = c + 1; r < MV_M; ++r) {
                const cd f = cd_div(A[r][c], A[c][c]);
                for (int j = c + 1; j < MV_M; ++j) A[r][j] = cd_sub(A[r][j], cd_mul(f, A[c][j]));
                for (int j = 0; j < MV_M; ++j) Bm[r][j] = cd_sub(Bm[r][j], cd_mul(f, Bm[c][j]));
            }
        }
        for (int j = 0; j < MV_M; ++j)
            for (int r = MV_M - 1; r >= 0; --r) {
                cd s = Bm[r][j];
                for (int c = r + 1; c < MV_M; ++c) s = cd_sub(s, cd_mul(A[r][c], Bm[c][j]));
                Bm[r][j] = cd_div(s, A[r][r]);
            }
        // back to complex64 (asteroid casts the solve result to the input dtype), trace, divide
        float tr_r = 0.f, tr_i = 0.f;
        for (int i = 0; i < MV_M; ++i) { tr_r += (float)Bm[i][i].x; tr_i += (float)Bm[i][i].y; }
        tr_r += 1e-8f;
        const float den = tr_r * tr_r + tr_i * tr_i;
        for (int i = 0; i < MV_M; ++i) {
            const float nr = (float)Bm[i][0].x, ni = (float)Bm[i][0].y;
            wv[mk][lane][i] = make_float2((nr * tr_r + ni * tr_i) / den, (ni * tr_r - nr * tr_i) / den);
        }
        // ---- phase 3: y[t] = sum_m conj(w_m) x_m[t] ----
        float2 w[MV_M];
#pragma unroll
        for (int i = 0; i < MV_M; ++i) w[i] = wv[mk][lane][i];
        float2* yrow = Y + ((b * 2 + mk) * (long)T) * CSS_KP + k;
        for (int t = 0; t < T; ++t) {
            float yr = 0.f, yi = 0.f;
#pragma unroll
            for (int d = 0; d < MV_M; ++d) {
                const float2 x = load_x(spec, edge, frames, n_chunks, b, hop_frames, d, t, k);
                yr += w[d].x * x.x + w[d].y * x.y;
                yi += w[d].x * x.y - w[d].y * x.x;
            }
            yrow[(long)t * CSS_KP] = make_float2(yr, yi);
            e_y += yr * yr + yi * yi;
        }
    }
    if (mk < 2) {   // warp-uniform: every lane of the two speech warps takes part
        e_y = warp_sum(e_y);
        e_mask = warp_sum(e_mask);
        if (lane == 0) { esum[mk][0] = e_y; esum[mk][1] = e_mask; }
    }
    __syncthreads();
    if (mk == 0 && lane < 4) {
        const int s = lane >> 1, w = lane & 1;
        energy[((b * 2 + s) * gridDim.x + blockIdx.x) * 2 + w] = esum[s][w];
    }
    // zero the bin padding of Y so later stages can read whole rows
    if (mk < 2 && !act && k < CSS_KP) {
        float2* yrow = Y + ((b * 2 + mk) * (long)T) * CSS_KP + k;
        for (int t = 0; t < T; ++t) yrow[(long)t * CSS_KP] = make_float2(0.f, 0.f);
    }
}

extern "C" int css_mvdr(const void* spec, const void* edge, long frames, long n_chunks, const float* stitched, long L,
                        int hop_frames, int T, void* Y, float* energy, cudaStream_t st) {
    CSS_REQUIRE(n_chunks > 0 && T > 1, "css_mvdr: empty input");
    CSS_REQUIRE(L >= (n_chunks - 1) * hop_frames + T, "css_mvdr: stitched masks too short (L=%ld)", L);
    CSS_REQUIRE(frames >= (n_chunks - 1) * hop_frames + T - 1, "css_mvdr: spectrogram too short (frames=%ld)", frames);
    CSS_REQUIRE(n_chunks <= 65535, "css_mvdr: more than 65535 chunks per call");
    dim3 grid((CSS_NBINS + MV_BINS - 1) / MV_BINS, (unsigned)n_chunks), block(MV_BINS, 3);
    mvdr_kernel<<<grid, block, 0, st>>>((const float2*)spec, (const float2*)edge, frames, n_chunks, stitched, L,
                                        hop_frames, T, (float2*)Y, energy);
    return css_check_launch("css_mvdr");
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
    chunk_scalars_kernel<<<(unsigned)((n_chunks + 127) / 128), 128, 0, st>>>(energy, n_chunks, nblk, T, rescale, scal);
    int rc = css_check_launch("chunk_scalars");
    if (rc) return rc;
    if (!rescale && !dedup) return 0;
    dim3 grid((T * CSS_KP + 255) / 256, (unsigned)n_chunks);
    mvdr_finalize_kernel<<<grid, 256, 0, st>>>((float2*)Y, scal, T, rescale, dedup);
    return css_check_launch("css_mvdr_finalize");
}
