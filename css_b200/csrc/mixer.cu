// On-the-fly training mixer on the GPU (reference: css/datasets/separation.py:163-267, which does this
// with lhotse cut arithmetic on the host: pad + mix two utterances at an offset, cut the mixture into windows, reverberate
// with probability 0.5, add noise at a sampled SNR with probability 0.5).  The utterance pool lives in HBM; the host only
// samples the descriptors (which utterances, which offset, which window), the kernels below do the arithmetic.
#include "common.cuh"

struct MixItem {          // one output window (mirrors css_mix_item in include/css_b200.h)
    long off1, len1;      // utterance 1 in the pool (per-channel sample offset / length); starts at mixture time 0
    long off2, len2;      // utterance 2, starts at mixture sample start2
    long start2;
    long win_start;       // first mixture sample of this window
};

// mix (B, C, W), s1 (B, W), s2 (B, W): sources are channel 0 of the padded utterances (separation.py:213-219,232-241)
__global__ void __launch_bounds__(256)
mix_windows_kernel(const float* __restrict__ pool, long ch_stride, int C, const MixItem* __restrict__ items, int W,
                   float* __restrict__ mix, float* __restrict__ s1, float* __restrict__ s2) {
    const int b = blockIdx.y;
    const MixItem it = items[b];
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < W; n += gridDim.x * blockDim.x) {
        const long t = it.win_start + n;
        const long t2 = t - it.start2;
        const bool in1 = t >= 0 && t < it.len1, in2 = t2 >= 0 && t2 < it.len2;
        for (int c = 0; c < C; ++c) {
            const float a = in1 ? pool[c * ch_stride + it.off1 + t] : 0.f;
            const float v = in2 ? pool[c * ch_stride + it.off2 + t2] : 0.f;
            mix[((long)b * C + c) * W + n] = a + v;
            if (c == 0) {
                s1[(long)b * W + n] = a;
                s2[(long)b * W + n] = v;
            }
        }
    }
}

extern "C" int css_mix_windows(const float* pool, long ch_stride, int C, const void* items, int B, int W, float* mix, float* s1,
                               float* s2, cudaStream_t st) {
    CSS_REQUIRE(pool && items && mix && s1 && s2 && B > 0 && W > 0 && C >= 1, "css_mix_windows: bad arguments");
    dim3 grid((W + 255) / 256 < 64 ? (W + 255) / 256 : 64, B);
    mix_windows_kernel<<<grid, 256, 0, st>>>(pool, ch_stride, C, static_cast<const MixItem*>(items), W, mix, s1, s2);
    return css_check_launch("css_mix_windows");
}

// CutMix at a target SNR (separation.py:150-157): x[b] += g[b] * noise[off[b] ...], g = sqrt(P_x / (P_n 10^(snr/10))) with the
// powers taken over channel 0 of the window; snr[b] = NaN leaves window b untouched (the transform fires with probability 0.5).
__global__ void __launch_bounds__(256)
add_noise_kernel(float* __restrict__ x, int C, int W, const float* __restrict__ noise, long noise_ch_stride, long noise_len,
                 const long* __restrict__ noise_off, const float* __restrict__ snr_db) {
    __shared__ double red[2][8];
    __shared__ float gain_s;
    const int b = blockIdx.x;
    const float snr = snr_db[b];
    if (isnan(snr)) return;
    const long off = noise_off[b];
    float* xb = x + (long)b * C * W;
    double px = 0.0, pn = 0.0;
    for (int n = threadIdx.x; n < W; n += blockDim.x) {
        const float a = xb[n], v = noise[(off + n) % noise_len];
        px += (double)(a * a);
        pn += (double)(v * v);
    }
    px = warp_sum_d(px);
    pn = warp_sum_d(pn);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = px; red[1][threadIdx.x >> 5] = pn; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double sx = 0.0, sn = 0.0;
        for (int i = 0; i < 8; ++i) { sx += red[0][i]; sn += red[1][i]; }
        gain_s = sn > 0.0 ? (float)sqrt(sx / (sn * pow(10.0, (double)snr / 10.0))) : 0.f;
    }
    __syncthreads();
    const float g = gain_s;
    for (int c = 0; c < C; ++c)
        for (int n = threadIdx.x; n < W; n += blockDim.x)
            xb[(long)c * W + n] += g * noise[c * noise_ch_stride + (off + n) % noise_len];
}

extern "C" int css_mix_add_noise(float* x, int B, int C, int W, const float* noise, long noise_ch_stride, long noise_len,
                                 const long* noise_off, const float* snr_db, cudaStream_t st) {
    CSS_REQUIRE(x && noise && noise_off && snr_db && B > 0 && W > 0 && noise_len > 0, "css_mix_add_noise: bad arguments");
    add_noise_kernel<<<B, 256, 0, st>>>(x, C, W, noise, noise_ch_stride, noise_len, noise_off, snr_db);
    return css_check_launch("css_mix_add_noise");
}

// ReverbWithImpulseResponse(normalize_output=True) (separation.py:140-148) as a direct FIR: y[n] = sum_k h[k] x[n-k] over the
// window (zero history), then y *= sqrt(E_x / E_y) so that the reverberated window keeps the energy of the dry one.
// rir_idx[b] < 0 leaves window b untouched.  One CTA per (window, channel); x is staged in shared memory tile by tile.
#define FIR_TILE 1024
__global__ void __launch_bounds__(256)
fir_kernel(const float* __restrict__ x, float* __restrict__ y, int W, const float* __restrict__ rirs, int taps,
           const int* __restrict__ rir_idx, int C, double* __restrict__ energy) {
    extern __shared__ float sm[];            // [taps] reversed h, then [FIR_TILE + taps] x
    const int b = blockIdx.x / C;
    const int r = rir_idx[b];
    if (r < 0) return;
    const float* h = rirs + (long)r * taps;
    float* hs = sm;
    float* xs = sm + taps;
    const float* xb = x + (long)blockIdx.x * W;
    float* yb = y + (long)blockIdx.x * W;
    for (int k = threadIdx.x; k < taps; k += blockDim.x) hs[k] = h[k];
    double ex = 0.0, ey = 0.0;
    for (int n0 = 0; n0 < W; n0 += FIR_TILE) {
        __syncthreads();
        for (int i = threadIdx.x; i < FIR_TILE + taps - 1; i += blockDim.x) {
            const int n = n0 - (taps - 1) + i;
            xs[i] = (n >= 0 && n < W) ? xb[n] : 0.f;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < FIR_TILE && n0 + i < W; i += blockDim.x) {
            float acc = 0.f;
            const float* xp = xs + i + taps - 1;         // x[n], then backwards
            for (int k = 0; k < taps; ++k) acc = fmaf(hs[k], xp[-k], acc);
            yb[n0 + i] = acc;
            const float xv = xp[0];
            ex += (double)(xv * xv);
            ey += (double)(acc * acc);
        }
    }
    ex = warp_sum_d(ex);
    ey = warp_sum_d(ey);
    if ((threadIdx.x & 31) == 0) { atomicAdd(energy + 2 * b, ex); atomicAdd(energy + 2 * b + 1, ey); }
}
__global__ void __launch_bounds__(256)
fir_normalize_kernel(float* __restrict__ y, int W, int C, const int* __restrict__ rir_idx, const double* __restrict__ energy) {
    const int b = blockIdx.y;
    if (rir_idx[b] < 0) return;
    const double ex = energy[2 * b], ey = energy[2 * b + 1];
    const float g = ey > 0.0 ? (float)sqrt(ex / ey) : 1.f;
    float* yb = y + (long)b * C * W;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < (long)C * W; i += (long)gridDim.x * blockDim.x) yb[i] *= g;
}

// x, y: (B, C, W), y != x; rirs (n_rirs, taps); rir_idx (B); energy: 2*B doubles of scratch.  Windows with rir_idx < 0 must
// be copied by the caller (y is only written for reverberated windows).
extern "C" int css_mix_reverb(const float* x, float* y, int B, int C, int W, const float* rirs, int taps, const int* rir_idx,
                              double* energy, cudaStream_t st) {
    CSS_REQUIRE(x && y && x != y && rirs && rir_idx && energy && B > 0 && W > 0, "css_mix_reverb: bad arguments");
    CSS_REQUIRE(taps >= 1 && taps <= 16384, "css_mix_reverb: %d taps unsupported (1..16384)", taps);
    const size_t smem = sizeof(float) * (size_t)(2 * taps + FIR_TILE);
    static CssPerDevice attr;
    if (attr.first()) cudaFuncSetAttribute(fir_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * (2 * 16384 + FIR_TILE));
    cudaMemsetAsync(energy, 0, sizeof(double) * 2 * B, st);
    fir_kernel<<<B * C, 256, smem, st>>>(x, y, W, rirs, taps, rir_idx, C, energy);
    int rc = css_check_launch("css_mix_reverb");
    if (rc) return rc;
    dim3 grid(32, B);
    fir_normalize_kernel<<<grid, 256, 0, st>>>(y, W, C, rir_idx, energy);
    return css_check_launch("css_mix_reverb (normalise)");
}
