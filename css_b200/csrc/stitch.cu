// K5-K7: DOA-based mask merge, inter-chunk permutation (overlap similarity), permutation-state
// scan, and winner-take-all overlap-average.
//   css/executor/separator.py:165-250, css/executor/stitcher.py:14-110.
// masks are (B, 257, T, C) fp32 (the reference layout), mags (B, 257, T).
#include "common.cuh"

// ---------------------------------------------------------------------------------------------
// block-wide sum of NV doubles per thread; result valid in thread 0 (and in red[] for all)
// ---------------------------------------------------------------------------------------------
template <int NV>
__device__ __forceinline__ void block_sum_d(double* v, double* red /*[32][NV]*/) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = warp_sum_d(v[i]);
    if (lane == 0)
        for (int i = 0; i < NV; ++i) red[warp * NV + i] = v[i];
    __syncthreads();
    if (warp == 0) {
        for (int i = 0; i < NV; ++i) {
            double s = (lane < nw) ? red[lane * NV + i] : 0.0;
            s = warp_sum_d(s);
            if (lane == 0) red[i] = s;
        }
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// K5: DOA likelihood + merge.  svT: (A, 7, nf) complex64 steering vectors of bins lo..lo+nf-1.
// ---------------------------------------------------------------------------------------------
#define DOA_A 30
#define DOA_M 7

__device__ __forceinline__ float pymod360(float x) {
    float r = fmodf(x, 360.f);
    return (r < 0.f) ? r + 360.f : r;
}

// PITCH: row pitch of the steering table in shared memory -- 64 when the band has at most 64 bins (the usual 80-2000 Hz band: 61), so
// that the 210 table reads of a point are one base register + immediates (with the runtime pitch nf every read carried its own
// address arithmetic: an eighth of the kernel's instructions); 0: pitch = nf.
template <int PITCH>
__global__ void __launch_bounds__(256, 2)      // two CTAs per SM: the 105 KB steering table of the usual band fits twice
doa_merge_kernel(const float2* __restrict__ spec, long stride_d, long stride_chunk, int T,
                 const float2* __restrict__ svT, const float* __restrict__ angles, int lo, int nf, float thresh,
                 float binarize, float* masks, int C, float* __restrict__ like_out, int* __restrict__ idx_out) {
    extern __shared__ float2 sv[];           // [A][7][nf]
    __shared__ double red[32 * 4];
    __shared__ double tot[2 * DOA_A];
    __shared__ int hit_s;
    const long b = blockIdx.x;
    const int pitch = PITCH ? PITCH : nf;
    for (int i = threadIdx.x; i < DOA_A * DOA_M * nf; i += blockDim.x) {
        const int row = i / nf;
        sv[row * pitch + (i - row * nf)] = svT[i];
    }
    __syncthreads();
    float acc[2][DOA_A];
#pragma unroll
    for (int a = 0; a < DOA_A; ++a) acc[0][a] = acc[1][a] = 0.f;
    float* mb = masks + b * (long)CSS_NBINS * T * C;
    const float2* sb = spec + (b * stride_chunk) * CSS_KP;
    for (int i = threadIdx.x; i < nf * T; i += blockDim.x) {
        const int t = i / nf, fi = i - t * nf, f = lo + fi;
        float2 xx[DOA_M], xy[DOA_M];         // (re, re) and (im, im) of each channel: operands of the packed FMAs below
        float xpow = 0.f;
#pragma unroll
        for (int m = 0; m < DOA_M; ++m) {
            const float2 x = sb[((long)m * stride_d + t) * CSS_KP + f];
            xpow += x.x * x.x + x.y * x.y;
            xx[m] = make_float2(x.x, x.x);
            xy[m] = make_float2(x.y, x.y);
        }
        const float* mk = mb + ((long)f * T + t) * C;
        const float b0 = mk[0] > binarize ? 1.f : 0.f, b1 = mk[1] > binarize ? 1.f : 0.f;
#pragma unroll
        for (int a = 0; a < DOA_A; ++a) {
            // conj(x) . s = (sum xr sr + sum xi si, sum xr si - sum xi sr): P = sum xr * (sr, si), Q = sum xi * (sr, si),
            // two packed fp32 FMAs (FFMA2) per channel -- the kernel is bound by the FMA pipe
            float2 P = make_float2(0.f, 0.f), Q = make_float2(0.f, 0.f);
#pragma unroll
            for (int m = 0; m < DOA_M; ++m) {
                const float2 sm = sv[(a * DOA_M + m) * pitch + fi];
                P = __ffma2_rn(xx[m], sm, P);
                Q = __ffma2_rn(xy[m], sm, Q);
            }
            const float re = P.x + Q.y, im = P.y - Q.x;
            // NaN when the radicand rounds < 0 (Q7): sqrt.approx keeps that, and its 1-2 ulp are below what the order of the
            // 9,150-term sums already differs from torch's by (IEEE sqrtf was six more instructions per angle and point)
            float l;
            asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(xpow - (re * re + im * im)));   // .ftz: no denormal scaling sequence around the MUFU
            l = -l;
            acc[0][a] += b0 * l;
            acc[1][a] += b1 * l;
        }
    }
    // reduce the 60 partials, 4 at a time (unrolled so acc[][] stays in registers)
#pragma unroll
    for (int a0 = 0; a0 < 2 * DOA_A; a0 += 4) {
        double v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int e = a0 + q;
            v[q] = (double)acc[e / DOA_A][e % DOA_A];
        }
        block_sum_d<4>(v, red);
        if (threadIdx.x < 4) tot[a0 + threadIdx.x] = red[threadIdx.x];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        int idx[2];
        for (int c = 0; c < 2; ++c) {
            int best = 0;
            bool nan_found = false;
            float bv = -INFINITY;
            for (int a = 0; a < DOA_A; ++a) {
                const float v = (float)tot[c * DOA_A + a];
                like_out[(b * 2 + c) * DOA_A + a] = v;
                if (nan_found) continue;
                if (isnan(v)) { best = a; nan_found = true; }
                else if (v > bv) { bv = v; best = a; }
            }
            idx[c] = best;
            idx_out[b * 3 + c] = best;
        }
        const float d0 = angles[idx[0]], d1 = angles[idx[1]];
        const int hit = fminf(pymod360(d0 - d1), pymod360(d1 - d0)) <= thresh;
        idx_out[b * 3 + 2] = hit;
        hit_s = hit;
    }
    __syncthreads();
    if (hit_s)   // victim is always output 0 (css/executor/separator.py:195-199)
        for (int i = threadIdx.x; i < CSS_NBINS * T; i += blockDim.x) mb[(long)i * C] = 1e-12f;
}

extern "C" int css_doa_merge(const void* spec, long stride_d, long stride_chunk, long n_chunks, int T, const void* svT,
                             const float* angles, int lo, int hi, float thresh, float* masks, int C, float* like,
                             int* idx, cudaStream_t st) {
    const int nf = hi - lo;
    CSS_REQUIRE(n_chunks > 0 && T > 0 && C >= 2, "css_doa_merge: bad shape");
    CSS_REQUIRE(lo >= 0 && hi <= CSS_NBINS && nf > 0, "css_doa_merge: bad band [%d,%d)", lo, hi);
    const int pitch = nf <= 64 ? 64 : nf;
    const size_t smem = sizeof(float2) * DOA_A * DOA_M * pitch;
    CSS_REQUIRE(smem <= 200 * 1024, "css_doa_merge: band of %d bins needs too much shared memory", nf);
    static CssPerDevice attr_set;
    if (attr_set.first()) {
        cudaFuncSetAttribute(doa_merge_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        cudaFuncSetAttribute(doa_merge_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    CssProfScope prof(7, (double)n_chunks * T * (CSS_NBINS * C * 4.0 + 7.0 * (hi - lo) * 8.0), st);
    if (pitch == 64)
        doa_merge_kernel<64><<<(unsigned)n_chunks, 256, smem, st>>>((const float2*)spec, stride_d, stride_chunk, T, (const float2*)svT,
                                                                    angles, lo, nf, thresh, 0.5f, masks, C, like, idx);
    else
        doa_merge_kernel<0><<<(unsigned)n_chunks, 256, smem, st>>>((const float2*)spec, stride_d, stride_chunk, T, (const float2*)svT,
                                                                   angles, lo, nf, thresh, 0.5f, masks, C, like, idx);
    return css_check_launch("css_doa_merge");
}

// ---------------------------------------------------------------------------------------------
// K6: permutation of adjacent chunks by overlap similarity (css/executor/stitcher.py:24-50).
// One CTA per pair; per-term arithmetic in fp32 as the reference, accumulation in fp64.
// swap[n] = 1 when the reference would emit [1,0]; sims[n] = {sim00, sim01, sim10, sim11} (fp32).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
stitch_perm_kernel(const float* __restrict__ mags, const float* __restrict__ masks, int T, int C, int margin,
                   int* __restrict__ swap, float* __restrict__ sims) {
    __shared__ double red[32 * 4];
    const long n = blockIdx.x;
    const float* mp = masks + n * (long)CSS_NBINS * T * C;
    const float* mn = mp + (long)CSS_NBINS * T * C;
    const float* gp = mags + n * (long)CSS_NBINS * T;
    const float* gn = gp + (long)CSS_NBINS * T;
    double s[4] = {0.0, 0.0, 0.0, 0.0};   // [i][j] -> i*2+j ; i indexes "now", j indexes "prev"
    for (int e = threadIdx.x; e < CSS_NBINS * margin; e += blockDim.x) {
        const int f = e / margin, tt = e - f * margin;
        const long ip = (long)f * T + (T - margin + tt), in = (long)f * T + tt;
        const float ap = fabsf(gp[ip]), an = fabsf(gn[in]);
        const float p0 = mp[ip * C] * ap, p1 = mp[ip * C + 1] * ap;
        const float n0 = mn[in * C] * an, n1 = mn[in * C + 1] * an;
        s[0] += (double)sqrtf(fabsf(p0 - n0));
        s[1] += (double)sqrtf(fabsf(p1 - n0));
        s[2] += (double)sqrtf(fabsf(p0 - n1));
        s[3] += (double)sqrtf(fabsf(p1 - n1));
    }
    block_sum_d<4>(s, red);
    if (threadIdx.x == 0) {
        const float s00 = -(float)red[0], s01 = -(float)red[1], s10 = -(float)red[2], s11 = -(float)red[3];
        const float keep = s00 + s11, sw = s01 + s10;
        swap[n] = (keep >= sw) ? 0 : 1;
        if (sims) { sims[n * 4] = s00; sims[n * 4 + 1] = s01; sims[n * 4 + 2] = s10; sims[n * 4 + 3] = s11; }
    }
}

extern "C" int css_stitch_perm(const float* mags, const float* masks, long n_chunks, int T, int C, int margin,
                               int* swap, float* sims, cudaStream_t st) {
    CSS_REQUIRE(n_chunks >= 1 && T > 0 && C >= 2, "css_stitch_perm: bad shape");
    CSS_REQUIRE(margin > 0 && margin <= T, "css_stitch_perm: margin %d out of range", margin);
    if (n_chunks == 1) return 0;   // no pairs: empty permutation list
    CssProfScope prof(8, (double)(n_chunks - 1) * 2.0 * 3.0 * CSS_NBINS * margin * 4.0, st);
    stitch_perm_kernel<<<(unsigned)(n_chunks - 1), 256, 0, st>>>(mags, masks, T, C, margin, swap, sims);
    return css_check_launch("css_stitch_perm");
}

// ---------------------------------------------------------------------------------------------
// Permutation state = prefix XOR of the swap bits (css/executor/stitcher.py:55-60).
// state[0] = 0, state[n+1] = state[n] ^ swap[n].  Single CTA, segmented scan.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
perm_state_kernel(const int* __restrict__ swap, long n_pairs, int* __restrict__ state) {
    __shared__ int par[1024];
    const long seg = (n_pairs + 1023) / 1024;
    const long s0 = threadIdx.x * seg, s1 = min(n_pairs, s0 + seg);
    int p = 0;
    for (long i = s0; i < s1; ++i) p ^= (swap[i] != 0);
    par[threadIdx.x] = p;
    __syncthreads();
    if (threadIdx.x == 0) {
        int run = 0;
        for (int i = 0; i < 1024; ++i) { int v = par[i]; par[i] = run; run ^= v; }
        state[0] = 0;
    }
    __syncthreads();
    int cur = par[threadIdx.x];
    for (long i = s0; i < s1; ++i) { cur ^= (swap[i] != 0); state[i + 1] = cur; }
}

extern "C" int css_perm_states(const int* swap, long n_chunks, int* state, cudaStream_t st) {
    CSS_REQUIRE(n_chunks >= 1, "css_perm_states: no chunks");
    perm_state_kernel<<<1, 1024, 0, st>>>(swap, n_chunks - 1, state);
    return css_check_launch("css_perm_states");
}

// ---------------------------------------------------------------------------------------------
// K7: reorder by state, winner-take-all (floor 1e-4), overlap-average at `hop` frames.
//   css/executor/stitcher.py:62-110.   out: [3][L][CSS_KP] (bin fastest), L = hop*(B-1)+T.
// 32x32 (frame, bin) tiles: reads coalesced along t, writes coalesced along bins.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
stitch_connect_kernel(const float* __restrict__ masks, const int* __restrict__ state, long B, int T, int C, int hop,
                      long L, float* __restrict__ out) {
    __shared__ float tile[3][32][33];
    const long g0 = (long)blockIdx.x * 32;
    const int f0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const long g = g0 + tx;
    for (int r = ty; r < 32; r += 8) {
        const int f = f0 + r;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, cnt = 0.f;
        if (g < L && f < CSS_NBINS) {
            long nlo = (g - T + hop) / hop;           // ceil((g - (T-1)) / hop)
            if (g - (T - 1) <= 0) nlo = 0;
            long nhi = g / hop;
            if (nhi > B - 1) nhi = B - 1;
            for (long n = nlo; n <= nhi; ++n) {
                const float* m = masks + ((n * CSS_NBINS + f) * (long)T + (g - n * hop)) * C;
                float m0 = m[0], m1 = m[1];
                const float m2 = m[2];
                if (state[n]) { const float t = m0; m0 = m1; m1 = t; }
                const float mx = fmaxf(fmaxf(m0, m1), m2);
                a0 += (m0 == mx) ? m0 : 1e-4f;
                a1 += (m1 == mx) ? m1 : 1e-4f;
                a2 += (m2 == mx) ? m2 : 1e-4f;
                cnt += 1.f;
            }
            if (cnt == 0.f) cnt = 1.f;
        } else {
            cnt = 1.f;
        }
        tile[0][tx][r] = a0 / cnt;
        tile[1][tx][r] = a1 / cnt;
        tile[2][tx][r] = a2 / cnt;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const long gg = g0 + r;
        const int f = f0 + tx;
        if (gg < L && f < CSS_KP) {
#pragma unroll
            for (int c = 0; c < 3; ++c) out[((long)c * L + gg) * CSS_KP + f] = (f < CSS_NBINS) ? tile[c][r][tx] : 0.f;
        }
    }
}

extern "C" int css_stitch_connect(const float* masks, const int* state, long n_chunks, int T, int C, int hop,
                                  float* out, cudaStream_t st) {
    CSS_REQUIRE(n_chunks >= 1 && T > 0 && C == 3 && hop > 0, "css_stitch_connect: bad shape (C must be 3)");
    const long L = (long)hop * (n_chunks - 1) + T;
    dim3 grid((unsigned)((L + 31) / 32), (CSS_KP + 31) / 32);
    CssProfScope prof(8, (double)n_chunks * CSS_NBINS * 3.0 * 4.0 * (T + hop), st);
    stitch_connect_kernel<<<grid, 256, 0, st>>>(masks, state, n_chunks, T, C, hop, L, out);
    return css_check_launch("css_stitch_connect");
}
