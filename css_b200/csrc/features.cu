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
    // mx is in [1e-20, ~2] here: one MUFU.RCP (div.approx adds a range-scaling sequence for denominators beyond 2^126)
    float rmx;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rmx) : "f"(mx));
    const float a = mx > 0.f ? mn * rmx : 0.f;
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

// ---------------------------------------------------------------------------------------------
// Fused front end of the mask network for the executor path (round 2): spectrum -> features (feature.py:105-134,156-175) ->
// the model's input MVN over time (conformer.py:95 / blstm.py:80) -> bf16 split planes for the embed GEMM, plus the magnitude
// (B, 257, T).  Replaces css_features + mvn_time_planes on that path: the fp32 feature matrix (832 MB per 10-min session, written
// once and read once) no longer exists, and every spectrum value is read ONCE per chunk (the two-kernel form read each
// 2x in features_kernel with no L2 reuse between the passes: 3.8 GB per session against 1.6 GB algorithmic).
//
// One CTA per (chunk, FP_BG = 8 bins): thread = (bin, frame slot), 32 frame slots, frames slot, slot + 32, ...  The unit
// cross-phasors of the chunk (n_pairs x T x 8 float2, 57.6 KB at T = 150) stay in shared memory, so each is formed once instead
// of twice; they are then overwritten in place by the IPD values, which are normalised over time from shared memory.
// Four block-wide reductions (shuffle over the 4 frame slots of a warp that share a bin, then the 8 warps through shared
// memory, fixed order -> deterministic): {sum mag, sum cos, sum sin}, {mag variance, sum IPD}, {sum mag feature, IPD
// variance}, {mag feature variance}.  257 = 32 x 8 + 1: the last bin group is shifted to bins 249..256 and recomputes seven
// bins to identical bits.
// ---------------------------------------------------------------------------------------------
#define FP_BG 8
#define FP_THREADS 256
#define FP_SLOTS (FP_THREADS / FP_BG)       // 32 frame slots
#define FP_WARPS (FP_THREADS / 32)
#define FP_NV 16                             // reduction slots: 1 + 2 * 6 used by round 1 (up to 1 + 2 * 7 fit)

// Sum over the threads of the CTA that share this thread's bin, v[0..N) in place (every thread gets the result), finished
// by ONE thread per (value, bin) into what the callers need -- bit i of inv_mask set: 1 / (sqrt(sum / (T-1)) + eps), the
// reciprocal of the unbiased standard deviation + eps; clear: sum / T, the mean -- so the divisions are not repeated by the 32
// threads of a bin.
template <int N>
__device__ __forceinline__ void fp_reduce(float (&v)[N], float* red, float* stat, int bin, int warp, int lane, int T, unsigned inv_mask) {
#pragma unroll
    for (int i = 0; i < N; ++i) {
        v[i] += __shfl_xor_sync(0xffffffffu, v[i], 8);
        v[i] += __shfl_xor_sync(0xffffffffu, v[i], 16);
    }
    if (lane < FP_BG) {
#pragma unroll
        for (int i = 0; i < N; ++i) red[(i * FP_WARPS + warp) * FP_BG + bin] = v[i];
    }
    __syncthreads();
    if (threadIdx.x < N * FP_BG) {
        const int i = threadIdx.x / FP_BG, b = threadIdx.x % FP_BG;
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < FP_WARPS; ++w) s += red[(i * FP_WARPS + w) * FP_BG + b];
        stat[i * FP_BG + b] = ((inv_mask >> i) & 1u) ? 1.f / (sqrtf(s / (T - 1)) + FEAT_EPS) : s / T;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = stat[i * FP_BG + bin];
}

// SAME_R: every pair has the same right channel ("1,0;2,0;..."), loaded once per frame -- the caller's claim (`shared_right`; the
// tables live in device memory and the host does not read them); the kernel checks it and traps on a false claim.  Without the claim
// the general variant runs (right and for any tables, 120 registers -> 2 CTAs per SM).
template <int NP, int PL, bool SAME_R>
__global__ void __launch_bounds__(FP_THREADS, SAME_R ? 3 : 2)
features_planes_kernel(const float2* __restrict__ spec, long stride_d, long stride_chunk, int T, const int* __restrict__ ipd_l,
                       const int* __restrict__ ipd_r, float* __restrict__ mag, __nv_bfloat16* __restrict__ planes,
                       long plane_stride, int ldp, float* __restrict__ feat, int ldf, int n_groups) {
    extern __shared__ float fp_smem[];
    float2* U = reinterpret_cast<float2*>(fp_smem);                  // [NP][T][FP_BG] phasors, later .x = IPD value
    float* MG = fp_smem + (size_t)2 * NP * T * FP_BG;                // [T][FP_BG] magnitude, later the mag feature
    float* red = MG + (size_t)T * FP_BG;                             // [FP_NV][FP_WARPS][FP_BG]
    float* stat = red + FP_NV * FP_WARPS * FP_BG;                    // [FP_NV][FP_BG]
    int* offs = reinterpret_cast<int*>(stat + FP_NV * FP_BG);        // [2][8] channel offsets of the pairs (elements < 2^31, checked by the host), [16] = shared right channel
    // bin groups fastest: the CTAs that share a chunk's 128-byte lines (spectrum rows, plane rows) run together
    const long b = blockIdx.x / n_groups;
    const int grp = blockIdx.x - (int)b * n_groups;
    const int k0 = min(grp * FP_BG, CSS_NBINS - FP_BG);
    const int bin = threadIdx.x & (FP_BG - 1), slot = threadIdx.x / FP_BG;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k = k0 + bin;
    const float2* base = spec + (b * stride_chunk) * CSS_KP + k;
    constexpr int NPA = NP > 0 ? NP : 1;

    constexpr bool same_r = SAME_R;
    constexpr int NR = SAME_R ? 1 : NPA;
    if (SAME_R) {
        bool sr = true;
        for (int p = 1; p < NP; ++p) sr &= ipd_r[p] == ipd_r[0];
        if (!sr) __trap();
    }
    if ((int)threadIdx.x < NP) {
        const long ol = (long)ipd_l[threadIdx.x] * stride_d * CSS_KP, orr = (long)ipd_r[threadIdx.x] * stride_d * CSS_KP;
        if ((ol | orr) >> 31) __trap();                 // 32-bit element offsets (the host checks 8 channels' worth)
        offs[threadIdx.x] = (int)ol;
        offs[8 + threadIdx.x] = (int)orr;
    }
    __syncthreads();
    int offl[NPA], offr[NR];
#pragma unroll
    for (int p = 0; p < NP; ++p) offl[p] = offs[p];
#pragma unroll
    for (int p = 0; p < (NP > 0 ? NR : 0); ++p) offr[p] = offs[8 + p];
    struct Frame { float2 x0, xl[NPA], xr[NR]; };
    auto fetch = [&](Frame& f, int t) {
        const float2* q = base + (long)t * CSS_KP;
        f.x0 = q[0];
#pragma unroll
        for (int p = 0; p < NP; ++p) f.xl[p] = q[offl[p]];
        if constexpr (NP > 0) {
#pragma unroll
            for (int p = 0; p < NR; ++p) f.xr[p] = q[offr[p]];
        }
    };

    // ---- round 1: magnitudes and unit cross-phasors into shared memory, their sums over time ----
    float acc[1 + 2 * NP];
#pragma unroll
    for (int i = 0; i < 1 + 2 * NP; ++i) acc[i] = 0.f;
    {
        Frame cur, nxt;
        if (slot < T) fetch(cur, slot);
        for (int t = slot; t < T; t += FP_SLOTS) {
            if (t + FP_SLOTS < T) fetch(nxt, t + FP_SLOTS);
            const float m = sqrtf(cur.x0.x * cur.x0.x + cur.x0.y * cur.x0.y);
            MG[t * FP_BG + bin] = m;
            acc[0] += fmaxf(m, FEAT_EPS);
            // common path for every pair, branch-free; the rare cases (purely real bins, vanishing / huge products: digital
            // silence) are flagged, contribute zero here and are redone below by the full unit_cross, one pair at a time in a
            // loop that is not unrolled (no calls, no spills, one copy of the slow code)
            unsigned rare = 0;
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                const float2 a = cur.xl[p], c = cur.xr[same_r ? 0 : p];
                const float2 w = cmulc(a, c);
                const float n2 = w.x * w.x + w.y * w.y;
                const bool odd = (a.y == 0.f && c.y == 0.f) || !(n2 >= 1e-30f && n2 <= 1e30f);
                rare |= (odd ? 1u : 0u) << p;
                float r;                                           // n2 is in [1e-30, 1e30] wherever the result is kept
                asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(n2));
                r = r * (1.5f - 0.5f * n2 * r * r);
                const float2 u = odd ? make_float2(0.f, 0.f) : make_float2(w.x * r, w.y * r);
                U[(p * T + t) * FP_BG + bin] = u;
                acc[1 + 2 * p] += u.x;
                acc[2 + 2 * p] += u.y;
            }
            if (rare) {
                const float2* q = base + (long)t * CSS_KP;
#pragma unroll 1
                for (int p = 0; p < NP; ++p)
                    if ((rare >> p) & 1u) U[(p * T + t) * FP_BG + bin] = unit_cross(q[offs[p]], q[offs[8 + p]]);
#pragma unroll
                for (int p = 0; p < NP; ++p)
                    if ((rare >> p) & 1u) {
                        const float2 u = U[(p * T + t) * FP_BG + bin];
                        acc[1 + 2 * p] += u.x;
                        acc[2 + 2 * p] += u.y;
                    }
            }
            cur = nxt;
        }
    }
    fp_reduce(acc, red, stat, bin, warp, lane, T, 0u);                 // acc = means over time
    const float m_mean = acc[0];
    // mag[b][bin][t]: consecutive threads walk t inside a bin row
    {
        float* mrow = mag + (b * CSS_NBINS + k0) * (long)T;
        for (int i = threadIdx.x; i < FP_BG * T; i += FP_THREADS) {
            const int bb = i / T, t = i - bb * T;
            mrow[i] = MG[t * FP_BG + bb];
        }
    }

    // ---- round 2: magnitude variance; IPD = atan2(sin - mean sin, cos - mean cos) in place of the phasors, their sums ----
    float* frow = feat ? feat + (b * T) * (long)ldf + k : nullptr;
    float a2[1 + NP];
#pragma unroll
    for (int i = 0; i < 1 + NP; ++i) a2[i] = 0.f;
    for (int t = slot; t < T; t += FP_SLOTS) {
        const float dm = fmaxf(MG[t * FP_BG + bin], FEAT_EPS) - m_mean;
        a2[0] = fmaf(dm, dm, a2[0]);
#pragma unroll
        for (int p = 0; p < NP; ++p) {
            float2* up = U + (p * T + t) * FP_BG + bin;
            const float2 u = *up;
            const float v = ipd_atan2(u.y - acc[2 + 2 * p], u.x - acc[1 + 2 * p]);
            up->x = v;
            if (frow) frow[(long)t * ldf + (p + 1) * CSS_NBINS] = v;
            a2[1 + p] += v;
        }
    }
    fp_reduce(a2, red, stat, bin, warp, lane, T, 1u);
    const float m_inv = a2[0];

    // ---- round 3: the magnitude feature (in place of the magnitude) and its sum; IPD variances ----
    float mean[1 + NP], a3[1 + NP];
#pragma unroll
    for (int p = 0; p < NP; ++p) mean[1 + p] = a2[1 + p];
#pragma unroll
    for (int i = 0; i < 1 + NP; ++i) a3[i] = 0.f;
    for (int t = slot; t < T; t += FP_SLOTS) {
        const float v0 = (fmaxf(MG[t * FP_BG + bin], FEAT_EPS) - m_mean) * m_inv;
        MG[t * FP_BG + bin] = v0;
        if (frow) frow[(long)t * ldf] = v0;
        a3[0] += v0;
#pragma unroll
        for (int p = 0; p < NP; ++p) {
            const float d = U[(p * T + t) * FP_BG + bin].x - mean[1 + p];
            a3[1 + p] = fmaf(d, d, a3[1 + p]);
        }
    }
    fp_reduce(a3, red, stat, bin, warp, lane, T, ~1u);
    mean[0] = a3[0];

    // ---- round 4: variance of the magnitude feature ----
    float a4[1] = {0.f};
    for (int t = slot; t < T; t += FP_SLOTS) {
        const float d = MG[t * FP_BG + bin] - mean[0];
        a4[0] = fmaf(d, d, a4[0]);
    }
    fp_reduce(a4, red, stat, bin, warp, lane, T, 1u);
    float inv[1 + NP];
    inv[0] = a4[0];
#pragma unroll
    for (int p = 0; p < NP; ++p) inv[1 + p] = a3[1 + p];

    // ---- emit the planes of (v - mean_T) / (std_T + eps) ----
    __nv_bfloat16* prow = planes + (b * T) * (long)ldp + k;
    for (int t = slot; t < T; t += FP_SLOTS) {
        __nv_bfloat16* o = prow + (long)t * ldp;
#pragma unroll
        for (int p = 0; p < 1 + NP; ++p) {
            const float v = p == 0 ? MG[t * FP_BG + bin] : U[((p > 0 ? p - 1 : 0) * T + t) * FP_BG + bin].x;
            float r = (v - mean[p]) * inv[p];
#pragma unroll
            for (int q = 0; q < PL; ++q) {
                const __nv_bfloat16 h = __float2bfloat16_rn(r);
                r -= __bfloat162float(h);
                o[q * plane_stride + p * CSS_NBINS] = h;
            }
        }
    }
    // zero the K padding of the rows (planes, and the fp32 features when they are written)
    if (grp == 0) {
        const int F = CSS_NBINS * (NP + 1), padp = ldp - F;
        for (int i = threadIdx.x; i < T * padp; i += FP_THREADS) {
            const int t = i / padp, c = i - t * padp;
            for (int q = 0; q < PL; ++q) planes[q * plane_stride + (b * T + t) * (long)ldp + F + c] = __float2bfloat16_rn(0.f);
        }
        if (feat) {
            const int padf = ldf - F;
            for (int i = threadIdx.x; i < T * padf; i += FP_THREADS) {
                const int t = i / padf, c = i - t * padf;
                feat[(b * T + t) * (long)ldf + F + c] = 0.f;
            }
        }
    }
}

static size_t fp_smem_bytes(int np, int T) {
    return sizeof(float) * ((size_t)2 * np * T * FP_BG + (size_t)T * FP_BG + FP_NV * FP_WARPS * FP_BG + FP_NV * FP_BG) + 17 * sizeof(int);
}
bool features_planes_supported(int T, int n_pairs) {
    return T > 1 && n_pairs >= 0 && n_pairs <= 7 && fp_smem_bytes(n_pairs, T) <= 200 * 1024;
}

template <int NP, int PL>
static int features_planes_go(const float2* spec, long stride_d, long stride_chunk, long n_chunks, int T, const int* ipd_l,
                              const int* ipd_r, bool shared_right, float* mag, void* planes, long plane_stride, int ldp,
                              float* feat, int ldf, cudaStream_t st) {
    const size_t smem = fp_smem_bytes(NP, T);
    static CssPerDevice attr_set;
    if (attr_set.first()) {
        cudaFuncSetAttribute(features_planes_kernel<NP, PL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        cudaFuncSetAttribute(features_planes_kernel<NP, PL, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    const int n_groups = (CSS_NBINS + FP_BG - 1) / FP_BG;
    CSS_REQUIRE(n_chunks * n_groups < (1L << 31), "css_features_planes: too many chunks (%ld)", n_chunks);
    const unsigned grid = (unsigned)(n_chunks * n_groups);
    __nv_bfloat16* pp = static_cast<__nv_bfloat16*>(planes);
    if (shared_right || NP <= 1)
        features_planes_kernel<NP, PL, true><<<grid, FP_THREADS, smem, st>>>(spec, stride_d, stride_chunk, T, ipd_l, ipd_r, mag, pp,
                                                                           plane_stride, ldp, feat, ldf, n_groups);
    else
        features_planes_kernel<NP, PL, false><<<grid, FP_THREADS, smem, st>>>(spec, stride_d, stride_chunk, T, ipd_l, ipd_r, mag, pp,
                                                                            plane_stride, ldp, feat, ldf, n_groups);
    return css_check_launch("css_features_planes");
}

// mag (B, 257, T); planes: n_planes x (B*T, ldp) bf16, ldp >= 257 (n_pairs + 1), padding zeroed; feat (B*T, ldf) fp32 or NULL.
int features_planes_launch(const void* spec, long stride_d, long stride_chunk, long n_chunks, int T, const int* ipd_l,
                           const int* ipd_r, int n_pairs, int shared_right, float* mag, void* planes, long plane_stride, int ldp, int n_planes,
                           float* feat, int ldf, cudaStream_t st) {
    CSS_REQUIRE(n_chunks > 0 && T > 1, "css_features_planes: empty input (chunks=%ld T=%d)", n_chunks, T);
    CSS_REQUIRE(features_planes_supported(T, n_pairs), "css_features_planes: T=%d with %d IPD pairs does not fit shared memory", T, n_pairs);
    CSS_REQUIRE(ldp >= CSS_NBINS * (n_pairs + 1) && (!feat || ldf >= CSS_NBINS * (n_pairs + 1)), "css_features_planes: pitch too small");
    CSS_REQUIRE(n_planes >= 1 && n_planes <= 3, "css_features_planes: 1..3 planes");
    CSS_REQUIRE(n_pairs == 0 || 8 * stride_d * CSS_KP < (1L << 31), "css_features_planes: channel stride %ld too large for 32-bit offsets", stride_d);
    // algorithmic bytes: spectrum read once per chunk, magnitude + planes written
    CssProfScope prof(6, (double)n_chunks * T * (CSS_NBINS * ((n_pairs ? 7 : 1) * 8.0 + 4.0) + (double)ldp * 2.0 * n_planes), st);
    const float2* s = static_cast<const float2*>(spec);
#define FP_ARGS s, stride_d, stride_chunk, n_chunks, T, ipd_l, ipd_r, shared_right != 0, mag, planes, plane_stride, ldp, feat, ldf, st
#define FP_CASE(NP)                                                          \
    case NP:                                                                 \
        if (n_planes == 1) return features_planes_go<NP, 1>(FP_ARGS);        \
        if (n_planes == 2) return features_planes_go<NP, 2>(FP_ARGS);        \
        return features_planes_go<NP, 3>(FP_ARGS)
    switch (n_pairs) {
        FP_CASE(0); FP_CASE(1); FP_CASE(2); FP_CASE(3); FP_CASE(4); FP_CASE(5); FP_CASE(6); FP_CASE(7);
    }
#undef FP_CASE
#undef FP_ARGS
    return -1;
}

// Stand-alone C-ABI form (tests, other callers): same outputs as css_features followed by the model-side MVN + plane split.
extern "C" int css_features_planes(const void* spec, long stride_d, long stride_chunk, long n_chunks, int T, const int* ipd_l,
                                   const int* ipd_r, int n_pairs, int shared_right, float* mag, void* planes, long plane_stride, int ldp,
                                   int n_planes, float* feat, int ldf, cudaStream_t st) {
    return features_planes_launch(spec, stride_d, stride_chunk, n_chunks, T, ipd_l, ipd_r, n_pairs, shared_right, mag, planes, plane_stride, ldp,
                                  n_planes, feat, ldf, st);
}
