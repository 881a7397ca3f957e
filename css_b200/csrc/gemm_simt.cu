// fp32 SIMT GEMM with fused epilogue:  C = R + alpha * act(A . W^T + bias)
//   A (M,K) row-major pitch lda, W (N,K) row-major pitch ldw (torch.nn.Linear layout), C (M,N) pitch ldc.
// This is the exact-fp32 path (and the checker for the tcgen05 path in gemm_tc.cu).  128x128x16
// tiles, 8x8 register micro-tiles, register-prefetch double buffering.
#include "common.cuh"
#include "gemm.h"

#define BM 128
#define BN 128
#define BK 16
#define LDS_PAD 4

__global__ void __launch_bounds__(256)
gemm_simt_kernel(const float* __restrict__ A, int lda, const float* __restrict__ W, int ldw, float* C,
                 int ldc, long M, int N, int K, GemmEpilogue ep) {
    __shared__ __align__(16) float As[2][BK][BM + LDS_PAD];
    __shared__ __align__(16) float Bs[2][BK][BN + LDS_PAD];
    const int tid = threadIdx.x;
    const long m0 = (long)blockIdx.x * BM;
    const int n0 = blockIdx.y * BN;
    const int lrow = tid >> 2, lk = (tid & 3) * 4;   // loader: rows lrow, lrow+64 ; k offset lk
    const int tx = tid & 15, ty = tid >> 4;

    float4 pa[2], pb[2];
    auto gload = [&](int k0) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            long m = m0 + lrow + 64 * i;
            pa[i] = (m < M) ? *reinterpret_cast<const float4*>(A + m * lda + k0 + lk) : make_float4(0, 0, 0, 0);
            int n = n0 + lrow + 64 * i;
            pb[i] = (n < N) ? *reinterpret_cast<const float4*>(W + (long)n * ldw + k0 + lk) : make_float4(0, 0, 0, 0);
        }
    };
    auto sstore = [&](int buf) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            int r = lrow + 64 * i;
            As[buf][lk + 0][r] = pa[i].x; As[buf][lk + 1][r] = pa[i].y;
            As[buf][lk + 2][r] = pa[i].z; As[buf][lk + 3][r] = pa[i].w;
            Bs[buf][lk + 0][r] = pb[i].x; Bs[buf][lk + 1][r] = pb[i].y;
            Bs[buf][lk + 2][r] = pb[i].z; Bs[buf][lk + 3][r] = pb[i].w;
        }
    };

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    gload(0);
    sstore(0);
    __syncthreads();
    const int nk = K / BK;
    for (int kt = 0; kt < nk; ++kt) {
        const int cur = kt & 1;
        if (kt + 1 < nk) gload((kt + 1) * BK);
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float4 a0 = *reinterpret_cast<const float4*>(&As[cur][k][ty * 4]);
            float4 a1 = *reinterpret_cast<const float4*>(&As[cur][k][64 + ty * 4]);
            float4 b0 = *reinterpret_cast<const float4*>(&Bs[cur][k][tx * 4]);
            float4 b1 = *reinterpret_cast<const float4*>(&Bs[cur][k][64 + tx * 4]);
            float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (kt + 1 < nk) {
            sstore(cur ^ 1);
            __syncthreads();
        }
    }

#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const long m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
            if (n >= N) continue;
            float v = acc[i][j];
            if (ep.bias) v += ep.bias[n];
            if (ep.act == 1) v = fmaxf(v, 0.f);
            if (ep.clamp_max > 0.f) v = fminf(v, ep.clamp_max);
            v *= ep.alpha;
            if (ep.residual) v += ep.residual[m * ep.ldr + n];
            C[m * ldc + n] = v;
        }
    }
}

int gemm_simt_launch(const float* A, int lda, const float* W, int ldw, float* C, int ldc, long M, int N, int K,
                     GemmEpilogue ep, cudaStream_t st) {
    CSS_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem M=%ld N=%d K=%d", M, N, K);
    CSS_REQUIRE(K % BK == 0, "gemm: K=%d must be a multiple of %d (pad with zeros)", K, BK);
    CSS_REQUIRE(lda % 4 == 0 && ldw % 4 == 0, "gemm: lda/ldw must be multiples of 4");
    CSS_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)W & 15) == 0, "gemm: A/W must be 16-byte aligned");
    dim3 grid((unsigned)((M + BM - 1) / BM), (N + BN - 1) / BN);
    gemm_simt_kernel<<<grid, 256, 0, st>>>(A, lda, W, ldw, C, ldc, M, N, K, ep);
    return css_check_launch("gemm_simt");
}

extern "C" int css_gemm_f32(const float* A, int lda, const float* W, int ldw, const float* bias, float* C, int ldc,
                            long M, int N, int K, int act, float alpha, const float* residual, int ldr,
                            float clamp_max, cudaStream_t st) {
    GemmEpilogue ep;
    ep.bias = bias; ep.act = act; ep.alpha = alpha; ep.residual = residual; ep.ldr = ldr; ep.clamp_max = clamp_max;
    return gemm_simt_launch(A, lda, W, ldw, C, ldc, M, N, K, ep, st);
}
