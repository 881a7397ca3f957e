// K2: fused spectral + spatial features (css/executor/feature.py:105-134,156-175) and the
// model-side input MVN over time (css/models/conformer.py:95, css/models/blstm.py:80).
#include "common.cuh"
#include <cuda_bf16.h>

#define FEAT_EPS 1.1920928955078125e-07f
#define FEAT_THREADS 96    // bins per CTA: three CTAs (3 warps each) cover the 257 bins of a chunk
#define FEAT_SPLIT 3        // 2,241 short CTAs for a 10-min session instead of 747 long ones in 1.7 waves
#define FEAT_TT 32         // time tile of the transposing mag store

// unit phasor of a * conj(b); atan2(0,0) = 0 convention -> (1, 0) when a product vanishes
__device__ __forceinline__ float2 unit_cross(float2 a, float2 b) {
    if (a.y == 0.f && b.y == 0.f) {
        // purely real bins (DC, Nyquist): the reference takes atan2(+0, re) in {0, pi_f}, so the phase
        // difference is 0 or +-pi_f and (cos, sin) = (-1, -+sin(pi_f)); the sign of that 8.7e-8 residue
        // decides between +pi and -pi in the final atan2 (feature.py:118-133) and must be reproduced.
        const bool na = signbit(a.x), nb = signbit(b.x);
        if (na == nb) return make_float2(1.f, 0.f);
        return make_float2(-1.f, na ? -8.742278e-8f : 8.742278e-8f);
    }
    float2 p = cmulc(a, b);
    const float n2 = p.x * p.x + p.y * p.y;
    if (n2 < 1e-30f || n2 > 1e30f) {
        // vanishing or huge product: normalise the operands first (phase(a) - phase(b), a zero operand's phase taken as 0)
        const float na = sqrtf(a.x * a.x + a.y * a.y), nb = sqrtf(b.x * b.x + b.y * b.y);
        if (na == 0.f && nb == 0.f) return make_float2(1.f, 0.f);
        if (na == 0.f) return make_float2(b.x / nb, -b.y / nb);
        if (nb == 0.f) return make_float2(a.x / na, a.y / na);
        a.x /= na; a.y /= na; b.x /= nb; b.y /= nb;
        return cmulc(a, b);
    }
    // one MUFU.RSQ + a Newton step (full fp32 accuracy) instead of sqrt and two IEEE divisions: this function runs
    // 2 x 6 x 150 times per bin and the kernel is bound by instruction issue, not by HBM
    float r = rsqrtf(n2);
    r = r * (1.5f - 0.5f * n2 * r * r);
    return make_float2(p.x * r, p.y * r);
}

// atan2f for the IPD (6 x 150 values per bin and chunk; the kernel is bound by instruction issue and libdevice's atan2f with its
// IEEE division and slow paths was a third of it): octant reduction, MUFU.RCP quotient, degree-8 polynomial in a^2 fitted at
// Chebyshev nodes (max error 1.1e-7 rad evaluated in fp32, 2e-6 with the quotient's 2 ulp; the parity gate on features is 1e-3).
// Signed zeros as atan2f: (+-0, x < 0 or -0) -> +-pi, (0, +0) -> 0 -- the DC/Nyquist choice of unit_cross relies on it.
__device__ __forceinline__ float ipd_atan2(float y, float x) {
    const float ax = fabsf(x), ay = fabsf(y);
    float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    if (mx < 1e-20f) { mx *= 1.8446744e19f; mn *= 1.8446744e19f; }      // keep the reciprocal in range
    const float a = mx > 0.f ? __fdividef(mn, mx) : 0.f;
    const float s = a * a;
    float r = 0.0028340641874819994f;
    r = fmaf(r, s, -0.016005029901862144f);
    r = fmaf(r, s, 0.042587608098983765f);
    r = fmaf(r, s, -0.07495445758104324f);
    r = fmaf(r, s, 0.10636754333972931f);
    r = fmaf(r, s, -0.14202570915222168f);
    r = fmaf(r, s, 0.19992484152317047f);
    r = fmaf(r, s, -0.3333306610584259f);
    r = fmaf(r * s, a, a);
    if (ay > ax) r = 1.57079637f - r;
    if (signbit(x)) r = 3.14159274f - r;
    return copysignf(r, y);
}

// FEAT_SPLIT blocks per chunk; a thread owns one bin.  spec element (d, chunk b, frame t, bin k) lives at
// spec[(d*stride_d + b*stride_chunk + t) * CSS_KP + k].
__global__ void __launch_bounds__(FEAT_THREADS, 9)
features_kernel(const float2* __restrict__ spec, long stride_d, long stride_chunk, int T,
                const int* __restrict__ ipd_l, const int* __restrict__ ipd_r, int n_pairs,
                float* __restrict__ mag, float* __restrict__ feat, int ldf) {
    __shared__ float tile[FEAT_TT][FEAT_THREADS + 1];
    const long b = blockIdx.x;
    const int k0 = blockIdx.y * FEAT_THREADS, k = k0 + threadIdx.x;
    const int nb = min(FEAT_THREADS, CSS_NBINS - k0);         // bins of this CTA
    const bool act = k < CSS_NBINS;
    const float2* base = spec + (b * stride_chunk) * CSS_KP + k;

    // per-pair channel pointers in registers (the index tables are read once, not once per frame), and the common case of
    // one shared reference channel ("1,0;2,0;...") loads that channel once per frame
    const float2* pl[8];
    const float2* pr[8];
    bool same_r = true;
#pragma unroll
    for (int p = 0; p < 8; ++p) {
        const int l = p < n_pairs ? ipd_l[p] : 0, r = p < n_pairs ? ipd_r[p] : 0;
        pl[p] = base + (long)l * stride_d * CSS_KP;
        pr[p] = base + (long)r * stride_d * CSS_KP;
        if (p < n_pairs && r != ipd_r[0]) same_r = false;
    }

    // ---- pass 1: statistics over time ----
    float m_sum = 0.f;
    float cs[8], sn[8];
#pragma unroll
    for (int p = 0; p < 8; ++p) cs[p] = sn[p] = 0.f;
    if (act) {
        for (int t = 0; t < T; ++t) {
            const long o = (long)t * CSS_KP;
            const float2 x0 = base[o];
            m_sum += fmaxf(sqrtf(x0.x * x0.x + x0.y * x0.y), FEAT_EPS);
            float2 xl[8], xr0 = make_float2(0.f, 0.f);
#pragma unroll
            for (int p = 0; p < 8; ++p)
                if (p < n_pairs) xl[p] = pl[p][o];
            if (same_r && n_pairs > 0) xr0 = pr[0][o];
#pragma unroll
            for (int p = 0; p < 8; ++p)
                if (p < n_pairs) {
                    const float2 u = unit_cross(xl[p], same_r ? xr0 : pr[p][o]);
                    cs[p] += u.x;
                    sn[p] += u.y;
                }
        }
    }
    const float m_mean = m_sum / T;
    float var = 0.f;
    if (act)
        for (int t = 0; t < T; ++t) {
            float2 x0 = base[(long)t * CSS_KP];
            float dlt = fmaxf(sqrtf(x0.x * x0.x + x0.y * x0.y), FEAT_EPS) - m_mean;
            var += dlt * dlt;
        }
    const float inv = 1.f / (sqrtf(var / (T - 1)) + FEAT_EPS);
#pragma unroll
    for (int p = 0; p < 8; ++p) { cs[p] /= T; sn[p] /= T; }

    // ---- pass 2: emit features (row-major (chunk, t) x feature) and transposed magnitude ----
    float* frow = feat + (b * T) * (long)ldf;
    float* mrow = mag + b * (long)CSS_NBINS * T;
    for (int t0 = 0; t0 < T; t0 += FEAT_TT) {
        const int tn = min(FEAT_TT, T - t0);
        if (act)
            for (int tt = 0; tt < tn; ++tt) {
                const int t = t0 + tt;
                const long o = (long)t * CSS_KP;
                const float2 x0 = base[o];
                float2 xl[8], xr0 = make_float2(0.f, 0.f);
#pragma unroll
                for (int p = 0; p < 8; ++p)
                    if (p < n_pairs) xl[p] = pl[p][o];
                if (same_r && n_pairs > 0) xr0 = pr[0][o];
                float m = sqrtf(x0.x * x0.x + x0.y * x0.y);
                tile[tt][threadIdx.x] = m;
                float* fr = frow + (long)t * ldf + k;
                fr[0] = (fmaxf(m, FEAT_EPS) - m_mean) * inv;
#pragma unroll
                for (int p = 0; p < 8; ++p)
                    if (p < n_pairs) {
                        const float2 u = unit_cross(xl[p], same_r ? xr0 : pr[p][o]);
                        fr[(p + 1) * CSS_NBINS] = ipd_atan2(u.y - sn[p], u.x - cs[p]);
                    }
            }
        __syncthreads();
        // mag[b][bin][t]: consecutive threads walk t inside a bin row
        for (int i = threadIdx.x; i < nb * tn; i += blockDim.x) {
            const int bin = i / tn, tt = i - bin * tn;
            mrow[(long)(k0 + bin) * T + t0 + tt] = tile[tt][bin];
        }
        __syncthreads();
    }
    // zero the K padding of the feature rows
    const int F = CSS_NBINS * (n_pairs + 1);
    if (blockIdx.y == 0)
    for (int i = threadIdx.x; i < T * (ldf - F); i += blockDim.x) {
        const int t = i / (ldf - F), c = i - t * (ldf - F);
        frow[(long)t * ldf + F + c] = 0.f;
    }
}

// In-place MVN over time of x[n][t][0..F) (row pitch ld): (x - mean_T) / (std_T,unbiased + eps).
__global__ void mvn_time_kernel(float* __restrict__ x, int T, int F, int ld) {
    const long n = blockIdx.x;
    const int f = blockIdx.y * blockDim.x + threadIdx.x;
    if (f >= F) return;
    float* col = x + n * (long)T * ld + f;
    float s = 0.f;
    for (int t = 0; t < T; ++t) s += col[(long)t * ld];
    const float mean = s / T;
    float v = 0.f;
    for (int t = 0; t < T; ++t) {
        float d = col[(long)t * ld] - mean;
        v += d * d;
    }
    const float inv = 1.f / (sqrtf(v / (T - 1)) + FEAT_EPS);
    for (int t = 0; t < T; ++t) col[(long)t * ld] = (col[(long)t * ld] - mean) * inv;
}

extern "C" int css_features(const void* spec, long stride_d, long stride_chunk, long n_chunks, int T,
                            const int* ipd_l, const int* ipd_r, int n_pairs, float* mag, float* feat, int ldf,
                            cudaStream_t st) {
    CSS_REQUIRE(n_chunks > 0 && T > 1, "css_features: empty input (chunks=%ld T=%d)", n_chunks, T);
    CSS_REQUIRE(n_pairs >= 0 && n_pairs <= 8, "css_features: at most 8 IPD pairs supported, got %d", n_pairs);
    CSS_REQUIRE(ldf >= CSS_NBINS * (n_pairs + 1), "css_features: feature pitch %d too small", ldf);
    CssProfScope prof(6, (double)n_chunks * T * CSS_NBINS * ((n_pairs ? 7 : 1) * 8.0 + (n_pairs + 1) * 4.0 + 4.0), st);
    features_kernel<<<dim3((unsigned)n_chunks, FEAT_SPLIT), FEAT_THREADS, 0, st>>>((const float2*)spec, stride_d, stride_chunk, T, ipd_l,
                                                                 ipd_r, n_pairs, mag, feat, ldf);
    return css_check_launch("css_features");
}

// Same normalisation, written as bf16 split planes (pitch ldp >= F, padding columns zeroed) for the tcgen05 embed GEMM;
// x is left untouched.  One CTA per (sequence, 64 features): the (T x 64) tile is staged in shared memory with all its
// loads in flight at once, then thread (feature, quarter of the frames) reduces and emits its share.
#define MVN_F 64
__global__ void __launch_bounds__(256)
mvn_time_planes_kernel(const float* __restrict__ x, int T, int F, int ld, __nv_bfloat16* __restrict__ planes,
                       long plane_stride, int ldp, int n_planes) {
    extern __shared__ float mvn_tile[];          // [T][MVN_F]
    __shared__ float red[4][MVN_F];
    __shared__ float stat[2][MVN_F];
    const long n = blockIdx.x;
    const int f0 = blockIdx.y * MVN_F;
    const int f = threadIdx.x & (MVN_F - 1), part = threadIdx.x >> 6;
    const float* src = x + n * (long)T * ld + f0 + f;
    const bool live = f0 + f < F;
    for (int t = part; t < T; t += 4) mvn_tile[t * MVN_F + f] = live ? src[(long)t * ld] : 0.f;
    __syncthreads();
    const int per = (T + 3) / 4;
    const int t0 = part * per, t1 = min(T, t0 + per);
    float s = 0.f;
    for (int t = t0; t < t1; ++t) s += mvn_tile[t * MVN_F + f];
    red[part][f] = s;
    __syncthreads();
    if (part == 0) stat[0][f] = (red[0][f] + red[1][f] + red[2][f] + red[3][f]) / T;
    __syncthreads();
    const float mean = stat[0][f];
    float v = 0.f;
    for (int t = t0; t < t1; ++t) {
        const float d = mvn_tile[t * MVN_F + f] - mean;
        v += d * d;
    }
    red[part][f] = v;
    __syncthreads();
    if (part == 0) stat[1][f] = 1.f / (sqrtf((red[0][f] + red[1][f] + red[2][f] + red[3][f]) / (T - 1)) + FEAT_EPS);
    __syncthreads();
    if (f0 + f >= ldp) return;
    const float inv = stat[1][f];
    __nv_bfloat16* out = planes + n * (long)T * ldp + f0 + f;
    for (int t = t0; t < t1; ++t) {
        float r = live ? (mvn_tile[t * MVN_F + f] - mean) * inv : 0.f;
        for (int p = 0; p < n_planes; ++p) {
            const __nv_bfloat16 h = __float2bfloat16_rn(r);
            r -= __bfloat162float(h);
            out[p * plane_stride + (long)t * ldp] = h;
        }
    }
}

int mvn_time_planes_launch(const float* x, long n, int T, int F, int ld, void* planes, long plane_stride, int ldp, int n_planes,
                           cudaStream_t st) {
    CSS_REQUIRE(n > 0 && T > 1 && F > 0 && ld >= F && ldp >= F, "mvn_time_planes: bad shape");
    CSS_REQUIRE(T <= 512, "mvn_time_planes: T=%d exceeds the shared-memory tile (512 frames)", T);
    const size_t smem = sizeof(float) * (size_t)T * MVN_F;
    static CssPerDevice attr_set;
    if (attr_set.first()) {
        cudaFuncSetAttribute(mvn_time_planes_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 512 * MVN_F * 4);
    }
    dim3 grid((unsigned)n, (ldp + MVN_F - 1) / MVN_F);
    CssProfScope prof(6, (double)n * T * (F * 4.0 + (double)ldp * 2.0 * n_planes), st);
    mvn_time_planes_kernel<<<grid, 256, smem, st>>>(x, T, F, ld, static_cast<__nv_bfloat16*>(planes), plane_stride, ldp, n_planes);
    return css_check_launch("mvn_time_planes");
}

extern "C" int css_mvn_time(float* x, long n, int T, int F, int ld, cudaStream_t st) {
    CSS_REQUIRE(n > 0 && T > 1 && F > 0 && ld >= F, "css_mvn_time: bad shape");
    dim3 grid((unsigned)n, (F + 127) / 128);
    mvn_time_kernel<<<grid, 128, 0, st>>>(x, T, F, ld);
    return css_check_launch("css_mvn_time");
}
