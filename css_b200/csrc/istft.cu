// K12: inverse STFT (torch.istft, center=True, Hann, window-envelope normalisation) fused with
// the slice placement of css/executor/beamformer.py:83-118, plus the 0.9/max|.| peak
// normalisation of beamformer.py:122-123.  Only the frames whose samples are consumed by the
// placement plan are transformed.
#include "fft.cuh"

#define IS_WARPS 9           // up to 8 output blocks of 256 samples need 9 frames
#define IS_BLOCKS 8

// plan item: {chunk, s_begin, s_end, dst}.  Samples [s_begin, s_end) of the chunk's istft output
// (s_end - s_begin <= 8*256, all inside blocks [s_begin/256, s_begin/256 + 8)) are ADDED to
// out[spk][dst ...].  Items of one launch must not overlap in the destination.
__global__ void __launch_bounds__(IS_WARPS * 32)
istft_place_kernel(const float2* __restrict__ Y, int T, const int4* __restrict__ plan, float* __restrict__ out,
                   long out_len) {
    __shared__ FftTables tab;
    __shared__ __align__(16) float2 buf[IS_WARPS][2][256];
    fft_tables_init(&tab);
    __syncthreads();
    const int4 it = plan[blockIdx.x];
    const int spk = blockIdx.y;
    const int j0 = it.y >> 8;                          // first output block
    const int nb = ((it.z - 1) >> 8) - j0 + 1;         // blocks touched (<= 8)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    FftLaneTw ltw;
    fft_lane_twiddles(ltw, tab.tw, lane);
    float* fw = reinterpret_cast<float*>(buf[warp][1]);   // windowed time-domain frame of this warp
    if (warp <= nb && j0 + warp < T) {
        const float2* X = Y + (((long)it.x * 2 + spk) * T + (j0 + warp)) * CSS_KP;
        float2* a = buf[warp][0];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int k = lane + 32 * i;
            float2 xk = X[k], xc = X[256 - k];
            if (k == 0) { xk.y = 0.f; xc.y = 0.f; }      // c2r ignores imag of DC / Nyquist
            a[k] = irfft_pack(xk, xc, tab.tw, k);
        }
        __syncwarp();
        const float2* z = warp_fft256(a, buf[warp][1], ltw, lane);   // == a after 4 passes
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int n = lane + 32 * i;
            const float2 v = z[n];
            fw[2 * n] = v.x * (1.f / 256.f) * tab.win[2 * n];
            fw[2 * n + 1] = -v.y * (1.f / 256.f) * tab.win[2 * n + 1];
        }
    }
    __syncthreads();
    float* dst = out + (long)spk * out_len;
    for (int e = threadIdx.x; e < nb * 256; e += blockDim.x) {
        const int q = e >> 8, i = e & 255;
        const int s = ((j0 + q) << 8) + i;
        if (s < it.y || s >= it.z) continue;
        const float* f0 = reinterpret_cast<const float*>(buf[q][1]);
        const float* f1 = reinterpret_cast<const float*>(buf[q + 1][1]);
        const float w0 = tab.win[256 + i], w1 = tab.win[i];
        const float v = (f0[256 + i] + f1[i]) / (w0 * w0 + w1 * w1);
        const long o = (long)it.w + (s - it.y);
        if (o >= 0 && o < out_len) dst[o] += v;
    }
}

extern "C" int css_istft_place(const void* Y, int T, const int* plan, int n_items, float* out, long out_len,
                               cudaStream_t st) {
    CSS_REQUIRE(T >= 2 && out_len > 0, "css_istft_place: bad shape");
    if (n_items == 0) return 0;
    dim3 grid((unsigned)n_items, 2);
    CssProfScope prof(10, (double)n_items * 8.0 * (CSS_NBINS * 8.0 + CSS_HOP * 4.0), st);   // <= 8 frames in, 8 x 256 samples out per item
    istft_place_kernel<<<grid, IS_WARPS * 32, 0, st>>>((const float2*)Y, T, (const int4*)plan, out, out_len);
    return css_check_launch("css_istft_place");
}

// ---------------------------------------------------------------------------------------------
// peak normalisation: x[r][:] *= 0.9 / max|x[r][:]|   (beamformer.py:122-123)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
absmax_kernel(const float* __restrict__ x, long n, long ld, unsigned* __restrict__ mx) {
    const float* row = x + (long)blockIdx.y * ld;
    float m = 0.f;
    bool nan = false;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const float v = fabsf(row[i]);
        nan |= isnan(v);
        m = fmaxf(m, v);
    }
    if (nan) m = __int_as_float(0x7fc00000);   // torch.max propagates NaN
    unsigned bits = __float_as_uint(m);        // order-preserving for non-negative floats; NaN sorts last
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) bits = max(bits, __shfl_xor_sync(0xffffffffu, bits, o));
    if ((threadIdx.x & 31) == 0) atomicMax(mx + blockIdx.y, bits);
}
__global__ void __launch_bounds__(256)
peak_scale_kernel(float* __restrict__ x, long n, long ld, const unsigned* __restrict__ mx, float target) {
    float* row = x + (long)blockIdx.y * ld;
    const float peak = __uint_as_float(mx[blockIdx.y]);
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
        row[i] = row[i] * target / peak;       // (x * 0.9) / max, the reference's order of operations
}

static dim3 peak_grid(int rows, long n) {
    dim3 grid((unsigned)((n + 256 * 8 - 1) / (256 * 8)), rows);
    if (grid.x > 148 * 8) grid.x = 148 * 8;
    if (grid.x < 1) grid.x = 1;
    return grid;
}

// The two halves of the peak normalisation as separate calls, for a recording whose samples are spread over several
// GPUs: every rank takes the maximum of its own sample range (bits[r] = max(bits[r], float bits of max|x[r][:n]|), bits must
// start at 0), the ranks combine the bits with an all-reduce(MAX) (order-preserving for non-negative floats), and
// every rank scales its own range.  x: (rows, ld) with n <= ld valid samples per row.
extern "C" int css_absmax(const float* x, int rows, long n, long ld, unsigned* bits, cudaStream_t st) {
    CSS_REQUIRE(rows > 0 && n >= 0 && ld >= n, "css_absmax: bad shape");
    if (n == 0) return 0;
    CssProfScope prof(10, (double)rows * n * 4.0, st);
    absmax_kernel<<<peak_grid(rows, n), 256, 0, st>>>(x, n, ld, bits);
    return css_check_launch("css_absmax");
}
extern "C" int css_peak_scale(float* x, int rows, long n, long ld, float target, const unsigned* bits, cudaStream_t st) {
    CSS_REQUIRE(rows > 0 && n >= 0 && ld >= n, "css_peak_scale: bad shape");
    if (n == 0) return 0;
    CssProfScope prof(10, (double)rows * n * 8.0, st);
    peak_scale_kernel<<<peak_grid(rows, n), 256, 0, st>>>(x, n, ld, bits, target);
    return css_check_launch("css_peak_scale");
}

extern "C" int css_peak_normalize(float* x, int rows, long n, float target, unsigned* scratch, cudaStream_t st) {
    CSS_REQUIRE(rows > 0 && n > 0, "css_peak_normalize: empty input");
    cudaMemsetAsync(scratch, 0, sizeof(unsigned) * rows, st);
    int rc = css_absmax(x, rows, n, n, scratch, st);
    if (rc) return rc;
    return css_peak_scale(x, rows, n, n, target, scratch, st);
}

// ---------------------------------------------------------------------------------------------
// css_peak_normalize without caller scratch and without any allocation (for callers that cannot allocate: CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  The per-row peak bits live in a file-scope device array
// that every GPU loading the library holds until exit: CSS_PEAK_NOALLOC_MAX_ROWS x 4 B = 256 B.  No counterpart in the
// reference.
//
// In host code `g_peak_bits` is only a stand-in for the symbol, not a device address: the address is looked up once per
// device (with that device current) by cudaGetSymbolAddress and that pointer is what the kernels get.
// All calls share the array and it keeps the last call's bits, so it is cleared at the start of every call (stream-ordered);
// calls that overlap on the GPU (different streams without an ordering between them) race on it.
// ---------------------------------------------------------------------------------------------
#define CSS_PEAK_NOALLOC_MAX_ROWS 64
static __device__ __align__(256) unsigned g_peak_bits[CSS_PEAK_NOALLOC_MAX_ROWS];

static int peak_bits_address(unsigned** out) {
    static unsigned* addr[CSS_MAX_DEVICES] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        css_set_error("css_peak_normalize_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
        return -2;
    }
    CSS_REQUIRE(dev >= 0 && dev < CSS_MAX_DEVICES, "css_peak_normalize_noalloc: device ordinal %d beyond %d", dev, CSS_MAX_DEVICES);
    if (!addr[dev]) {
        unsigned* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_peak_bits);
        if (e != cudaSuccess || !p) {
            css_set_error("css_peak_normalize_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return -2;
        }
        addr[dev] = p;
    }
    *out = addr[dev];
    return 0;
}

extern "C" int css_peak_normalize_noalloc_max_rows(void) { return CSS_PEAK_NOALLOC_MAX_ROWS; }

extern "C" int css_peak_normalize_noalloc(float* x, int rows, long n, float target, cudaStream_t st) {
    CSS_REQUIRE(rows > 0 && n > 0, "css_peak_normalize_noalloc: empty input");
    CSS_REQUIRE(rows <= CSS_PEAK_NOALLOC_MAX_ROWS, "css_peak_normalize_noalloc: %d rows exceed the built-in scratch of %d rows",
                rows, CSS_PEAK_NOALLOC_MAX_ROWS);
    unsigned* bits = nullptr;
    int rc = peak_bits_address(&bits);
    if (rc) return rc;
    cudaError_t e = cudaMemsetAsync(bits, 0, sizeof(unsigned) * rows, st);
    if (e != cudaSuccess) {
        css_set_error("css_peak_normalize_noalloc: cudaMemsetAsync: %s", cudaGetErrorString(e));
        return -2;
    }
    // launched directly (not through css_absmax / css_peak_scale): their optional profiling scope creates events
    absmax_kernel<<<peak_grid(rows, n), 256, 0, st>>>(x, n, n, bits);
    rc = css_check_launch("css_peak_normalize_noalloc(absmax)");
    if (rc) return rc;
    peak_scale_kernel<<<peak_grid(rows, n), 256, 0, st>>>(x, n, n, bits, target);
    return css_check_launch("css_peak_normalize_noalloc(scale)");
}
