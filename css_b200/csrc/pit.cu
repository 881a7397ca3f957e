// K13: utterance-level PIT-MSE (css/objectives/pit.py:5-62, css/objectives/mse.py:29-37).
// One pass builds the K x K pairwise squared-error matrix of an utterance; a warp then searches
// the K! permutations in itertools order (lane-parallel, first minimum wins) and the gradient
// is written for the selected permutation only.
#include "common.cuh"

#define PIT_MAXK 6

// decode the idx-th permutation of range(K) in lexicographic (itertools.permutations) order
__device__ __forceinline__ void nth_perm(int idx, int K, int* perm) {
    int fact[PIT_MAXK + 1];
    fact[0] = 1;
    for (int i = 1; i <= K; ++i) fact[i] = fact[i - 1] * i;
    unsigned used = 0;
    for (int i = 0; i < K; ++i) {
        const int f = fact[K - 1 - i];
        int r = idx / f;
        idx -= r * f;
        for (int c = 0; c < K; ++c) {
            if (used & (1u << c)) continue;
            if (r == 0) { perm[i] = c; used |= 1u << c; break; }
            --r;
        }
    }
}

// est, tgt: (B, K, n) contiguous, n = T*F.   pair[b][i][j] = sum (est[b][i] - tgt[b][j])^2
__global__ void __launch_bounds__(256)
pit_pairwise_kernel(const float* __restrict__ est, const float* __restrict__ tgt, int K, long n,
                    double* __restrict__ pair) {
    __shared__ double red[32];
    const long b = blockIdx.y;
    const int ij = blockIdx.x, i = ij / K, j = ij % K;
    const float* e = est + (b * K + i) * n;
    const float* t = tgt + (b * K + j) * n;
    double acc = 0.0;
    for (long x = threadIdx.x; x < n; x += blockDim.x) {
        const float d = e[x] - t[x];
        acc += (double)(d * d);
    }
    acc = warp_sum_d(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double v = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.0;
        v = warp_sum_d(v);
        if (threadIdx.x == 0) pair[(b * K + i) * K + j] = v;
    }
}

// torch.min ordering over the stacked candidate losses (css/objectives/pit.py:52-56): NaN is the minimum and the first
// NaN wins; otherwise the smaller loss, first index on ties.
__device__ __forceinline__ bool pit_better(float l, int p, float best, int best_idx) {
    if (isnan(l)) return !isnan(best) || p < best_idx;
    if (isnan(best)) return false;
    return l < best || (l == best && p < best_idx);
}

// one warp per utterance
__global__ void __launch_bounds__(32)
pit_search_kernel(const double* __restrict__ pair, int K, long n, float* __restrict__ loss, int* __restrict__ perm_out) {
    const long b = blockIdx.x;
    const int lane = threadIdx.x;
    int nperm = 1;
    for (int i = 2; i <= K; ++i) nperm *= i;
    float best = INFINITY;
    int best_idx = 0x7fffffff;
    for (int p = lane; p < nperm; p += 32) {
        int perm[PIT_MAXK];
        nth_perm(p, K, perm);
        double s = 0.0;
        for (int k = 0; k < K; ++k) s += pair[(b * K + perm[k]) * K + k];   // estimate[perm[k]] vs target[k]
        const float l = (float)(s / (double)(K * n));
        if (pit_better(l, p, best, best_idx)) { best = l; best_idx = p; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, best_idx, o);
        if (pit_better(ob, oi, best, best_idx)) { best = ob; best_idx = oi; }
    }
    if (lane == 0) {
        int perm[PIT_MAXK];
        if (best_idx < 0 || best_idx >= nperm) best_idx = 0;     // unreachable (every lane < nperm competes); keeps perm[] defined
        nth_perm(best_idx, K, perm);
        loss[b] = best;
        for (int k = 0; k < K; ++k) perm_out[b * K + k] = perm[k];
    }
}

__global__ void pit_mean_kernel(const float* __restrict__ loss, long B, float* __restrict__ out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        float s = 0.f;
        for (long b = 0; b < B; ++b) s += loss[b];    // same left-to-right order as mse.py:32-37
        out[0] = s / (float)B;
    }
}

// grad[b][perm[k]] = gscale * 2/(K n B) * (est[b][perm[k]] - tgt[b][k])
__global__ void __launch_bounds__(256)
pit_grad_kernel(const float* __restrict__ est, const float* __restrict__ tgt, const int* __restrict__ perm, int K,
                long n, long B, float gscale, float* __restrict__ grad) {
    const long b = blockIdx.z;
    const int k = blockIdx.y;
    const int src = perm[b * K + k];
    const float c = gscale * 2.f / ((float)K * (float)n * (float)B);
    const float* e = est + (b * K + src) * n;
    const float* t = tgt + (b * K + k) * n;
    float* g = grad + (b * K + src) * n;
    for (long x = (long)blockIdx.x * blockDim.x + threadIdx.x; x < n; x += (long)gridDim.x * blockDim.x)
        g[x] = c * (e[x] - t[x]);
}

extern "C" int css_pit_mse(const float* est, const float* tgt, long B, int K, long n, double* pair_ws, float* loss_b,
                           int* perm, float* loss, float* grad, float gscale, cudaStream_t st) {
    CSS_REQUIRE(B > 0 && n > 0, "css_pit_mse: empty input");
    CSS_REQUIRE(K >= 1 && K <= PIT_MAXK, "css_pit_mse: %d sources unsupported (1..%d)", K, PIT_MAXK);
    CSS_REQUIRE(B <= 65535, "css_pit_mse: batch too large");
    dim3 g1(K * K, (unsigned)B);
    pit_pairwise_kernel<<<g1, 256, 0, st>>>(est, tgt, K, n, pair_ws);
    int rc = css_check_launch("pit_pairwise");
    if (rc) return rc;
    pit_search_kernel<<<(unsigned)B, 32, 0, st>>>(pair_ws, K, n, loss_b, perm);
    if ((rc = css_check_launch("pit_search"))) return rc;
    pit_mean_kernel<<<1, 32, 0, st>>>(loss_b, B, loss);
    if ((rc = css_check_launch("pit_mean"))) return rc;
    if (grad) {
        dim3 g2((unsigned)((n + 256 * 4 - 1) / (256 * 4)), K, (unsigned)B);
        pit_grad_kernel<<<g2, 256, 0, st>>>(est, tgt, perm, K, n, B, gscale, grad);
        rc = css_check_launch("pit_grad");
    }
    return rc;
}
