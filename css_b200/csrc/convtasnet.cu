// Conv-TasNet building blocks (reference css/models/conv_tasnet.py:17-330): everything of the network that
// is not a 1x1 convolution.  Activations are kept TIME-MAJOR, (n * T, C) row-major, so every Conv1d(kernel 1) of the
// reference is one css_gemm_f32 call (A = activations, W = the conv weight (C_out, C_in)), and the kernels below cover the
// rest: the strided encoder, channel / global layer norms with PReLU, the dilated depthwise conv and the transposed-conv
// decoder with its overlap-add.
#include "common.cuh"

// ---- encoder: Conv1d(1, N, L, stride) (conv_tasnet.py:76): out[(n, t)][c] = b[c] + sum_k x[n][stride t + k] w[c][k] ----
__global__ void __launch_bounds__(256)
ctn_encoder_kernel(const float* __restrict__ x, long len, const float* __restrict__ w, const float* __restrict__ b, int N, int L,
                   int stride, long T, float* __restrict__ out) {
    extern __shared__ float seg[];            // [L]
    const long row = blockIdx.x;              // n * T + t
    const long n = row / T, t = row - n * T;
    for (int k = threadIdx.x; k < L; k += blockDim.x) seg[k] = x[n * len + t * stride + k];
    __syncthreads();
    for (int c = threadIdx.x; c < N; c += blockDim.x) {
        float a = b ? b[c] : 0.f;
        for (int k = 0; k < L; ++k) a = fmaf(seg[k], w[c * L + k], a);
        out[row * N + c] = a;
    }
}
extern "C" int css_ctn_encoder(const float* x, long n, long len, const float* w, const float* b, int N, int L, int stride, float* out,
                               cudaStream_t st) {
    CSS_REQUIRE(x && w && out && n > 0 && len >= L && N > 0 && L > 0 && stride > 0, "css_ctn_encoder: bad arguments");
    const long T = (len - L) / stride + 1;
    CSS_REQUIRE(n * T < (1L << 31), "css_ctn_encoder: too many frames");
    ctn_encoder_kernel<<<(unsigned)(n * T), 256, L * sizeof(float), st>>>(x, len, w, b, N, L, stride, T, out);
    return css_check_launch("css_ctn_encoder");
}

// ---- channel layer norm (CumulativeLayerNorm = nn.LayerNorm over channels per frame, conv_tasnet.py:205-225): warp per row ----
__global__ void __launch_bounds__(256)
row_layernorm_kernel(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ b, long M, int C, float eps,
                     float* __restrict__ y) {
    const long row = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= M) return;
    const int lane = threadIdx.x & 31;
    const float* xr = x + row * C;
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += xr[c];
    const float mean = warp_sum(s) / C;
    float q = 0.f;
    for (int c = lane; c < C; c += 32) { const float d = xr[c] - mean; q += d * d; }
    const float rstd = rsqrtf(warp_sum(q) / C + eps);
    for (int c = lane; c < C; c += 32) y[row * C + c] = (xr[c] - mean) * rstd * (g ? g[c] : 1.f) + (b ? b[c] : 0.f);
}
extern "C" int css_row_layernorm(const float* x, const float* g, const float* b, long M, int C, float eps, float* y, cudaStream_t st) {
    CSS_REQUIRE(x && y && M > 0 && C > 0, "css_row_layernorm: bad arguments");
    row_layernorm_kernel<<<(unsigned)((M + 7) / 8), 256, 0, st>>>(x, g, b, M, C, eps, y);
    return css_check_launch("css_row_layernorm");
}

// ---- PReLU + global layer norm (GlobalLayerNorm over (C, T) per sample, conv_tasnet.py:168-202) ----
// stats[n] = {sum, sum of squares} of prelu(x) over the T x C elements of sample n (fp64)
__global__ void __launch_bounds__(256)
ctn_prelu_stats_kernel(const float* __restrict__ x, long per_sample, const float* __restrict__ alpha, double* __restrict__ stats) {
    __shared__ double red[2][8];
    const long n = blockIdx.y;
    const float a = alpha[0];
    const float* xs = x + n * per_sample;
    double s = 0.0, q = 0.0;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < per_sample; i += (long)gridDim.x * blockDim.x) {
        float v = xs[i];
        v = v > 0.f ? v : a * v;
        s += (double)v;
        q += (double)v * (double)v;
    }
    s = warp_sum_d(s);
    q = warp_sum_d(q);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = s; red[1][threadIdx.x >> 5] = q; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double ss = 0.0, qq = 0.0;
        for (int i = 0; i < 8; ++i) { ss += red[0][i]; qq += red[1][i]; }
        atomicAdd(stats + 2 * n, ss);
        atomicAdd(stats + 2 * n + 1, qq);
    }
}
extern "C" int css_ctn_prelu_stats(const float* x, long n, long per_sample, const float* alpha, double* stats, cudaStream_t st) {
    CSS_REQUIRE(x && alpha && stats && n > 0 && per_sample > 0 && n <= 65535, "css_ctn_prelu_stats: bad arguments");
    cudaMemsetAsync(stats, 0, sizeof(double) * 2 * n, st);
    dim3 grid(148, (unsigned)n);
    ctn_prelu_stats_kernel<<<grid, 256, 0, st>>>(x, per_sample, alpha, stats);
    return css_check_launch("css_ctn_prelu_stats");
}

__device__ __forceinline__ float ctn_norm(float v, float a, float mean, float rstd, float g, float b) {
    v = v > 0.f ? v : a * v;
    return (v - mean) * rstd * g + b;
}
// y = gLN(prelu(x)) with the statistics above: y[(n,t)][c] = g[c] (prelu(x) - mean_n) / sqrt(var_n + eps) + b[c]
__global__ void __launch_bounds__(256)
ctn_norm_apply_kernel(const float* __restrict__ x, long T, int C, const float* __restrict__ alpha, const double* __restrict__ stats,
                      const float* __restrict__ g, const float* __restrict__ b, float eps, float* __restrict__ y) {
    const long n = blockIdx.y;
    const double cnt = (double)T * C;
    const double mean = stats[2 * n] / cnt;
    const float mu = (float)mean, rstd = rsqrtf((float)(stats[2 * n + 1] / cnt - mean * mean) + eps), a = alpha[0];
    const long base = n * T * C;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < T * C; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        y[base + i] = ctn_norm(x[base + i], a, mu, rstd, g[c], b[c]);
    }
}
extern "C" int css_ctn_norm_apply(const float* x, long n, long T, int C, const float* alpha, const double* stats, const float* g,
                                  const float* b, float eps, float* y, cudaStream_t st) {
    CSS_REQUIRE(x && y && alpha && stats && g && b && n > 0 && T > 0 && C > 0 && n <= 65535, "css_ctn_norm_apply: bad arguments");
    dim3 grid(148 * 2, (unsigned)n);
    ctn_norm_apply_kernel<<<grid, 256, 0, st>>>(x, T, C, alpha, stats, g, b, eps, y);
    return css_check_launch("css_ctn_norm_apply");
}

// ---- depthwise dilated conv over time on gLN(prelu(x)) (Conv1D_Block.dwconv, conv_tasnet.py:283-291,311): the input
// normalisation is applied on the fly, zero padding applies to the NORMALISED signal (non-causal: pad = dilation (P-1)/2) ----
__global__ void __launch_bounds__(256)
ctn_dwconv_kernel(const float* __restrict__ x, long T, int C, int P, int dil, const float* __restrict__ alpha,
                  const double* __restrict__ stats, const float* __restrict__ g, const float* __restrict__ b, float eps,
                  const float* __restrict__ w, const float* __restrict__ wb, float* __restrict__ y) {
    const long n = blockIdx.y;
    const double cnt = (double)T * C;
    const double mean = stats[2 * n] / cnt;
    const float mu = (float)mean, rstd = rsqrtf((float)(stats[2 * n + 1] / cnt - mean * mean) + eps), a = alpha[0];
    const long base = n * T * C;
    const int pad = dil * (P - 1) / 2;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < T * C; i += (long)gridDim.x * blockDim.x) {
        const long t = i / C;
        const int c = (int)(i - t * C);
        float acc = wb ? wb[c] : 0.f;
        const float gc = g[c], bc = b[c];
        for (int k = 0; k < P; ++k) {
            const long tt = t - pad + (long)k * dil;
            if (tt >= 0 && tt < T) acc = fmaf(w[c * P + k], ctn_norm(x[base + tt * C + c], a, mu, rstd, gc, bc), acc);
        }
        y[base + i] = acc;
    }
}
extern "C" int css_ctn_dwconv(const float* x, long n, long T, int C, int P, int dil, const float* alpha, const double* stats,
                              const float* g, const float* b, float eps, const float* w, const float* wb, float* y, cudaStream_t st) {
    CSS_REQUIRE(x && y && x != y && alpha && stats && g && b && w && n > 0 && T > 0 && C > 0 && P >= 1 && (P & 1) && dil >= 1 && n <= 65535,
                "css_ctn_dwconv: bad arguments (odd kernel size, non-causal)");
    dim3 grid(148 * 2, (unsigned)n);
    ctn_dwconv_kernel<<<grid, 256, 0, st>>>(x, T, C, P, dil, alpha, stats, g, b, eps, w, wb, y);
    return css_check_launch("css_ctn_dwconv");
}

// ---- decoder: d_s = w * relu-mask_s, ConvTranspose1d(N, 1, L, stride) (conv_tasnet.py:84,148-156):
//      frames[(s, n, t)][k] = sum_c enc[(n,t)][c] masks[(n,t)][s N + c] wd[c][k];  y[s][n][stride t + k] += frames + bias ----
__global__ void __launch_bounds__(128)
ctn_decode_frames_kernel(const float* __restrict__ enc, const float* __restrict__ masks, int ldm, int N, int L, int S, long rows,
                         const float* __restrict__ wd, float* __restrict__ frames) {
    extern __shared__ float d[];               // [N] masked encoder row
    __shared__ float red[4][32];
    const long row = blockIdx.x;
    const int s = blockIdx.y;
    for (int c = threadIdx.x; c < N; c += blockDim.x) d[c] = enc[row * N + c] * masks[row * ldm + s * N + c];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int k = 0; k < L; ++k) {
        float a = 0.f;
        for (int c = threadIdx.x; c < N; c += blockDim.x) a = fmaf(d[c], wd[c * L + k], a);
        a = warp_sum(a);
        if (lane == 0) red[warp][k & 31] = a;
        __syncthreads();
        if (threadIdx.x == 0) frames[((long)s * rows + row) * L + k] = red[0][k & 31] + red[1][k & 31] + red[2][k & 31] + red[3][k & 31];
        __syncthreads();
    }
}
__global__ void __launch_bounds__(256)
ctn_overlap_add_kernel(const float* __restrict__ frames, long T, int L, int stride, long out_len, const float* __restrict__ bias,
                       float* __restrict__ y) {
    const long sn = blockIdx.y;                // s * n_batch + n
    const float* f = frames + sn * T * L;
    for (long j = (long)blockIdx.x * blockDim.x + threadIdx.x; j < out_len; j += (long)gridDim.x * blockDim.x) {
        float a = bias ? bias[0] : 0.f;
        const long t_hi = j / stride;
        for (long t = t_hi; t >= 0 && j - t * stride < L; --t)
            if (t < T) a += f[t * L + (j - t * stride)];
        y[sn * out_len + j] = a;
    }
}
// enc (n T, N), masks (n T, ldm >= S N) -> y (S, n, (T-1) stride + L); frames: S * n * T * L floats of scratch
extern "C" int css_ctn_decode(const float* enc, const float* masks, int ldm, long n, long T, int N, int L, int stride, int S,
                              const float* wd, const float* bias, float* frames, float* y, cudaStream_t st) {
    CSS_REQUIRE(enc && masks && wd && frames && y && n > 0 && T > 0 && N > 0 && L > 0 && L <= 32 && S >= 1 && ldm >= S * N,
                "css_ctn_decode: bad arguments (filter length <= 32)");
    CSS_REQUIRE(n * T < (1L << 31) && S * n <= 65535, "css_ctn_decode: too many frames");
    dim3 g1((unsigned)(n * T), S);
    ctn_decode_frames_kernel<<<g1, 128, N * sizeof(float), st>>>(enc, masks, ldm, N, L, S, n * T, wd, frames);
    int rc = css_check_launch("css_ctn_decode (frames)");
    if (rc) return rc;
    const long out_len = (T - 1) * stride + L;
    dim3 g2(64, (unsigned)(S * n));
    ctn_overlap_add_kernel<<<g2, 256, 0, st>>>(frames, T, L, stride, out_len, bias, y);
    return css_check_launch("css_ctn_decode (overlap-add)");
}
