// K1/K8: batched windowed framing + shared-memory FFT-512 (forward STFT).
//
// Replaces css/executor/feature.py:45-82 (conv1d-as-DFT, uncentred) and the per-chunk
// centred torch.stft of css/executor/beamformer.py:137-145.  Layout: spec[sig][frame][CSS_KP]
// (complex64, bin fastest) so every consumer reads coalesced along bins.
#include "fft.cuh"

#define STFT_WARPS 8

// frame g of signal s covers samples [256 g, 256 g + 512).
__global__ void __launch_bounds__(STFT_WARPS * 32)
stft_kernel(const float* __restrict__ x, long n_sig, long sig_stride, long n_frames, float2* __restrict__ spec, long spec_stride) {
    __shared__ FftTables tab;
    __shared__ __align__(16) float2 buf[STFT_WARPS][2][256];
    fft_tables_init(&tab);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    FftLaneTw ltw;
    fft_lane_twiddles(ltw, tab.tw, lane);
    const long total = n_sig * n_frames;
    for (long w = (long)blockIdx.x * STFT_WARPS + warp; w < total; w += (long)gridDim.x * STFT_WARPS) {
        const long s = w / n_frames, g = w - s * n_frames;
        const float* src = x + s * sig_stride + g * CSS_HOP;
        float2* a = buf[warp][0];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int n = lane + 32 * i;
            a[n] = make_float2(__ldg(src + 2 * n) * tab.win[2 * n], __ldg(src + 2 * n + 1) * tab.win[2 * n + 1]);
        }
        __syncwarp();
        const float2* Z = warp_fft256(a, buf[warp][1], ltw, lane);
        float2* dst = spec + (s * spec_stride + g) * CSS_KP;
        for (int k = lane; k < CSS_KP; k += 32)
            dst[k] = (k < CSS_NBINS) ? rfft_bin(Z, tab.tw, k) : make_float2(0.f, 0.f);
        __syncwarp();
    }
}

// Centred frame 0 of each chunk: reflect-padded [x[256..1], x[0..255]] at chunk start
// (torch.stft(center=True, pad_mode="reflect")).  edge[sig][chunk][CSS_KP].
__global__ void __launch_bounds__(STFT_WARPS * 32)
stft_edge_kernel(const float* __restrict__ x, long n_sig, long sig_stride, long n_chunks, long chunk_hop,
                 float2* __restrict__ edge) {
    __shared__ FftTables tab;
    __shared__ __align__(16) float2 buf[STFT_WARPS][2][256];
    fft_tables_init(&tab);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    FftLaneTw ltw;
    fft_lane_twiddles(ltw, tab.tw, lane);
    const long total = n_sig * n_chunks;
    for (long w = (long)blockIdx.x * STFT_WARPS + warp; w < total; w += (long)gridDim.x * STFT_WARPS) {
        const long s = w / n_chunks, c = w - s * n_chunks;
        const float* src = x + s * sig_stride + c * chunk_hop;
        float2* a = buf[warp][0];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int n = lane + 32 * i;
            const int p0 = 2 * n, p1 = 2 * n + 1;   // padded positions; sample index = |p - 256|
            const int i0 = p0 < 256 ? 256 - p0 : p0 - 256;
            const int i1 = p1 < 256 ? 256 - p1 : p1 - 256;
            a[n] = make_float2(__ldg(src + i0) * tab.win[p0], __ldg(src + i1) * tab.win[p1]);
        }
        __syncwarp();
        const float2* Z = warp_fft256(a, buf[warp][1], ltw, lane);
        float2* dst = edge + (s * n_chunks + c) * CSS_KP;
        for (int k = lane; k < CSS_KP; k += 32)
            dst[k] = (k < CSS_NBINS) ? rfft_bin(Z, tab.tw, k) : make_float2(0.f, 0.f);
        __syncwarp();
    }
}

extern "C" int css_stft(const float* x, long n_sig, long sig_stride, long n_frames, void* spec, cudaStream_t st) {
    CSS_REQUIRE(n_sig > 0 && n_frames > 0, "css_stft: empty input (n_sig=%ld n_frames=%ld)", n_sig, n_frames);
    CSS_REQUIRE(sig_stride >= (n_frames - 1) * CSS_HOP + CSS_NFFT, "css_stft: signal stride %ld too short for %ld frames",
                sig_stride, n_frames);
    long warps = n_sig * n_frames;
    int grid = (int)((warps + STFT_WARPS - 1) / STFT_WARPS);
    if (grid > 148 * 8) grid = 148 * 8;
    CssProfScope prof(5, (double)warps * (CSS_HOP * 4 + CSS_NBINS * 8), st);
    stft_kernel<<<grid, STFT_WARPS * 32, 0, st>>>(x, n_sig, sig_stride, n_frames, (float2*)spec, n_frames);
    return css_check_launch("css_stft");
}

// Frames [0, n_frames) of signals whose first sample is x + s * sig_stride, written to spec + s * spec_stride frames: lets a
// caller fill a frame range of a larger [sig][frames][KP] buffer (recordings streamed to the device in segments).
extern "C" int css_stft_range(const float* x, long n_sig, long sig_stride, long n_frames, void* spec, long spec_stride,
                              cudaStream_t st) {
    CSS_REQUIRE(n_sig > 0 && n_frames > 0 && spec_stride >= n_frames, "css_stft_range: bad shape (n_sig=%ld n_frames=%ld)", n_sig, n_frames);
    long warps = n_sig * n_frames;
    int grid = (int)((warps + STFT_WARPS - 1) / STFT_WARPS);
    if (grid > 148 * 8) grid = 148 * 8;
    CssProfScope prof(5, (double)warps * (CSS_HOP * 4 + CSS_NBINS * 8), st);
    stft_kernel<<<grid, STFT_WARPS * 32, 0, st>>>(x, n_sig, sig_stride, n_frames, (float2*)spec, spec_stride);
    return css_check_launch("css_stft_range");
}

extern "C" int css_stft_edge(const float* x, long n_sig, long sig_stride, long n_chunks, long chunk_hop, void* edge,
                             cudaStream_t st) {
    CSS_REQUIRE(n_sig > 0 && n_chunks > 0, "css_stft_edge: empty input");
    CSS_REQUIRE(sig_stride >= (n_chunks - 1) * chunk_hop + 257, "css_stft_edge: signal stride too short");
    long warps = n_sig * n_chunks;
    int grid = (int)((warps + STFT_WARPS - 1) / STFT_WARPS);
    if (grid > 148 * 8) grid = 148 * 8;
    stft_edge_kernel<<<grid, STFT_WARPS * 32, 0, st>>>(x, n_sig, sig_stride, n_chunks, chunk_hop, (float2*)edge);
    return css_check_launch("css_stft_edge");
}
