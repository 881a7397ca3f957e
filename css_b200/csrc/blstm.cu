// BLSTM mask estimator forward (css/models/blstm.py:68-88,122-128,156-160).
// Per layer: one GEMM projects all T inputs of both directions (x W_ih^T + b_ih + b_hh), then T
// recurrence steps each run ONE kernel for both directions: a (chunks x H) . (H x 4H) product
// tiled 64 x (32 hidden units x 4 gates) with the LSTM cell fused in the epilogue, so the gates
// never touch HBM.  Chunks are the batch axis that keeps the step kernel wide.
// Two step kernels: fp32 FMAs (GEMM mode 0) and, in the tensor-core modes, mma.sync bf16 with the three-term split
// (h and W_hh kept as bf16 hi/lo planes; the epilogue writes the planes of h_t for the next step).
#include "common.cuh"
#include <cuda_bf16.h>
#include "conformer.h"
#include "gemm.h"
#include "../../include/css_b200.h"
#include <cstdlib>
#include <new>
#include <vector>

int gemm_dispatch(const float* A, int lda, const float* W, int ldw, float* C, int ldc, long M, int N, int K,
                  GemmEpilogue ep, cudaStream_t st);

#define LS_BN 64     // sequences per CTA
#define LS_BJ 32     // hidden units per CTA (x4 gates = 128 columns)
#define LS_BK 16

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// xproj: (N, T, 8H) = [fwd i f g o | bwd i f g o];  w_hh[dir]: (4H, H);  h_prev/h_next: (2, N, H);
// c: (2, N, H);  out: (N, T, 2H) = [fwd h | bwd h].   step s: fwd t = s, bwd t = T-1-s.
__global__ void __launch_bounds__(256)
lstm_step_kernel(const float* __restrict__ xproj, const float* __restrict__ w_hh_f, const float* __restrict__ w_hh_b,
                 const float* __restrict__ h_prev, float* __restrict__ h_next, float* __restrict__ c_state,
                 float* __restrict__ out, long N, int T, int H, int s) {
    __shared__ __align__(16) float Hs[2][LS_BK][LS_BN + 4];
    __shared__ __align__(16) float Ws[2][LS_BK][4 * LS_BJ + 4];
    const int dir = blockIdx.z;
    const int t = dir ? (T - 1 - s) : s;
    const float* W = dir ? w_hh_b : w_hh_f;
    const float* hp = h_prev + (long)dir * N * H;
    const long n0 = (long)blockIdx.x * LS_BN;
    const int j0 = blockIdx.y * LS_BJ;
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;       // thread: rows ty*4..+3, hidden units 2tx, 2tx+1
    const int lk = (tid & 3) * 4;
    const int hrow = tid >> 2;                    // 0..63 : sequence row for the H tile
    float4 ph, pw[2];
    auto gload = [&](int k0) {
        const long n = n0 + hrow;
        ph = (n < N) ? *reinterpret_cast<const float4*>(hp + n * H + k0 + lk) : make_float4(0, 0, 0, 0);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int col = hrow + 64 * i;        // 0..127 = gate*32 + jj
            const int row = (col >> 5) * H + j0 + (col & 31);
            pw[i] = *reinterpret_cast<const float4*>(W + (long)row * H + k0 + lk);
        }
    };
    auto sstore = [&](int buf) {
        Hs[buf][lk + 0][hrow] = ph.x; Hs[buf][lk + 1][hrow] = ph.y;
        Hs[buf][lk + 2][hrow] = ph.z; Hs[buf][lk + 3][hrow] = ph.w;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int col = hrow + 64 * i;
            Ws[buf][lk + 0][col] = pw[i].x; Ws[buf][lk + 1][col] = pw[i].y;
            Ws[buf][lk + 2][col] = pw[i].z; Ws[buf][lk + 3][col] = pw[i].w;
        }
    };
    float acc[4][4][2];   // [row][gate][unit]
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int g = 0; g < 4; ++g) acc[r][g][0] = acc[r][g][1] = 0.f;
    gload(0);
    sstore(0);
    __syncthreads();
    const int nk = H / LS_BK;
    for (int kt = 0; kt < nk; ++kt) {
        const int cur = kt & 1;
        if (kt + 1 < nk) gload((kt + 1) * LS_BK);
#pragma unroll
        for (int k = 0; k < LS_BK; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(&Hs[cur][k][ty * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const float2 w = *reinterpret_cast<const float2*>(&Ws[cur][k][g * LS_BJ + 2 * tx]);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    acc[r][g][0] = fmaf(av[r], w.x, acc[r][g][0]);
                    acc[r][g][1] = fmaf(av[r], w.y, acc[r][g][1]);
                }
            }
        }
        if (kt + 1 < nk) {
            sstore(cur ^ 1);
            __syncthreads();
        }
    }
    // fused cell update (gate order i, f, g, o as torch.nn.LSTM)
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const long n = n0 + ty * 4 + r;
        if (n >= N) continue;
        const float* xp = xproj + (n * T + t) * (long)(8 * H) + (long)dir * 4 * H;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int j = j0 + 2 * tx + u;
            const float gi = acc[r][0][u] + xp[j];
            const float gf = acc[r][1][u] + xp[H + j];
            const float gg = acc[r][2][u] + xp[2 * H + j];
            const float go = acc[r][3][u] + xp[3 * H + j];
            const long si = ((long)dir * N + n) * H + j;
            const float c = sigmoidf_(gf) * c_state[si] + sigmoidf_(gi) * tanhf(gg);
            const float h = sigmoidf_(go) * tanhf(c);
            c_state[si] = c;
            h_next[si] = h;
            out[(n * T + t) * (long)(2 * H) + (long)dir * H + j] = h;
        }
    }
}


// ---------------------------------------------------------------------------------------------
// Tensor-core step (GEMM modes 1, 2): gates = h_{t-1} . W_hh^T as mma.sync.m16n8k16 bf16 over hi/lo planes,
// hi.hi + hi.lo + lo.hi in fp32 accumulators (the same split as the tcgen05 GEMMs and the attention kernel).
// Warp tile: 32 sequences x (8 hidden units x 4 gates), so a thread's accumulators hold i, f, g, o of the same
// (sequence, unit) and the cell update needs no exchange.  CTA = WM x WN warps; K = H streamed in 64-wide chunks
// through a 3-stage cp.async ring.  whh planes: [dir][plane][4H][H] bf16; h planes: [dir][plane][N][H] bf16.
// ---------------------------------------------------------------------------------------------
namespace {
constexpr int LM_BK = 64;
constexpr int LM_ROWB = 144;          // bytes per smem row: 64 bf16 + 16 B pad (conflict-free ldmatrix)
constexpr int LM_STAGES = 3;

__device__ __forceinline__ void lm_mma(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void lm_ldsm4(uint32_t addr, uint32_t* r) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void lm_cp16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
// ex2.approx-based forms (abs error ~1e-7): the libm expf/tanhf of the fp32 step kernel were a third of this kernel's instructions
__device__ __forceinline__ float lm_sigmoid(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float lm_tanh(float x) {
    const float e = __expf(-2.f * fabsf(x));
    return copysignf(__fdividef(1.f - e, 1.f + e), x);
}
struct LmCellIn { float2 xi, xf, xg, xo, c; };
// gate pre-activations of (sequence n, time t, units j, j+1) and the cell state: requested before the k loop, used after it
__device__ __forceinline__ LmCellIn lm_cell_load(const float* __restrict__ xproj, const float* __restrict__ c_state, long n, long N, int T,
                                                 int t, int H, int dir, int j) {
    LmCellIn in;
    if (n >= N) n = N - 1;
    const float* xp = xproj + (n * T + t) * (long)(8 * H) + (long)dir * 4 * H + j;
    in.xi = __ldg(reinterpret_cast<const float2*>(xp));
    in.xf = __ldg(reinterpret_cast<const float2*>(xp + H));
    in.xg = __ldg(reinterpret_cast<const float2*>(xp + 2 * H));
    in.xo = __ldg(reinterpret_cast<const float2*>(xp + 3 * H));
    in.c = *reinterpret_cast<const float2*>(c_state + ((long)dir * N + n) * H + j);
    return in;
}
// a0..a3: accumulators of gates i, f, g, o for units j, j+1 (the [2*half], [2*half+1] entries of an m16n8 tile)
__device__ __forceinline__ void lm_cell_store(const LmCellIn& in, float ai0, float ai1, float af0, float af1, float ag0, float ag1, float ao0,
                                              float ao1, float* __restrict__ c_state, float* __restrict__ out, __nv_bfloat16* __restrict__ hn,
                                              long plane_h, long n, long N, int T, int t, int H, int dir, int j) {
    float2 c, h;
    c.x = lm_sigmoid(af0 + in.xf.x) * in.c.x + lm_sigmoid(ai0 + in.xi.x) * lm_tanh(ag0 + in.xg.x);
    c.y = lm_sigmoid(af1 + in.xf.y) * in.c.y + lm_sigmoid(ai1 + in.xi.y) * lm_tanh(ag1 + in.xg.y);
    h.x = lm_sigmoid(ao0 + in.xo.x) * lm_tanh(c.x);
    h.y = lm_sigmoid(ao1 + in.xo.y) * lm_tanh(c.y);
    *reinterpret_cast<float2*>(c_state + ((long)dir * N + n) * H + j) = c;
    *reinterpret_cast<float2*>(out + (n * T + t) * (long)(2 * H) + (long)dir * H + j) = h;
    const __nv_bfloat162 hi = __floats2bfloat162_rn(h.x, h.y);
    const __nv_bfloat162 lo = __floats2bfloat162_rn(h.x - __bfloat162float(hi.x), h.y - __bfloat162float(hi.y));
    *reinterpret_cast<__nv_bfloat162*>(hn + n * H + j) = hi;
    *reinterpret_cast<__nv_bfloat162*>(hn + plane_h + n * H + j) = lo;
}

template <int WM, int WN>
struct LmCfg {
    static constexpr int BM = 32 * WM;            // sequences per CTA
    static constexpr int JB = 8 * WN;             // hidden units per CTA
    static constexpr int BROWS = 4 * JB;          // W_hh rows per CTA: [gate][unit]
    static constexpr int A_PLANE = BM * LM_ROWB;
    static constexpr int B_PLANE = BROWS * LM_ROWB;
    static constexpr int STAGE = 2 * A_PLANE + 2 * B_PLANE;
    static constexpr int SMEM = LM_STAGES * STAGE;
    static constexpr int THREADS = 32 * WM * WN;
};

template <int WM, int WN>
__global__ void __launch_bounds__(32 * WM * WN)
lstm_step_mma_kernel(const float* __restrict__ xproj, const __nv_bfloat16* __restrict__ whh, const __nv_bfloat16* __restrict__ h_prev,
                     __nv_bfloat16* __restrict__ h_next, float* __restrict__ c_state, float* __restrict__ out, long N, int T, int H, int s) {
    using C = LmCfg<WM, WN>;
    extern __shared__ __align__(128) uint8_t lm_smem[];
    const int dir = blockIdx.z;
    const int t = dir ? (T - 1 - s) : s;
    const long n0 = (long)blockIdx.x * C::BM;
    const int j0 = blockIdx.y * C::JB;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wm = warp / WN, wn = warp - wm * WN;
    const long plane_h = N * (long)H, plane_w = 4L * H * H;
    const __nv_bfloat16* hp = h_prev + (long)dir * 2 * plane_h;
    const __nv_bfloat16* wp = whh + (long)dir * 2 * plane_w;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(lm_smem);

    // every thread copies the same (row, chunk) slots of each stage: sources and destinations are worked out once
    constexpr int A_IT = 2 * C::BM * 8 / C::THREADS, B_IT = 2 * C::BROWS * 8 / C::THREADS;
    static_assert(A_IT * C::THREADS == 2 * C::BM * 8 && B_IT * C::THREADS == 2 * C::BROWS * 8, "copy slots must divide evenly");
    const __nv_bfloat16* asrc[A_IT];
    const __nv_bfloat16* bsrc[B_IT];
    uint32_t adst[A_IT], bdst[B_IT];
#pragma unroll
    for (int it = 0; it < A_IT; ++it) {
        const int rs = (tid >> 3) + it * (C::THREADS / 8), p = rs / C::BM, r = rs % C::BM;
        long n = n0 + r;
        if (n >= N) n = N - 1;                                         // rows past N are computed and dropped
        asrc[it] = hp + p * plane_h + n * H + (tid & 7) * 8;
        adst[it] = sbase + p * C::A_PLANE + r * LM_ROWB + (tid & 7) * 16;
    }
#pragma unroll
    for (int it = 0; it < B_IT; ++it) {
        const int rs = (tid >> 3) + it * (C::THREADS / 8), p = rs / C::BROWS, r = rs % C::BROWS;
        const int g = r / C::JB, u = r - g * C::JB;
        bsrc[it] = wp + p * plane_w + ((long)g * H + j0 + u) * H + (tid & 7) * 8;
        bdst[it] = sbase + 2 * C::A_PLANE + p * C::B_PLANE + r * LM_ROWB + (tid & 7) * 16;
    }
    auto load_stage = [&](int kc, int stage) {
#pragma unroll
        for (int it = 0; it < A_IT; ++it) lm_cp16(adst[it] + stage * C::STAGE, asrc[it] + kc * LM_BK);
#pragma unroll
        for (int it = 0; it < B_IT; ++it) lm_cp16(bdst[it] + stage * C::STAGE, bsrc[it] + kc * LM_BK);
    };

    float acc[2][4][4];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int g = 0; g < 4; ++g)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[m][g][i] = 0.f;

    const int gid = lane >> 2, q = lane & 3;
    const int j = j0 + wn * 8 + 2 * q;
    const int nk = H / LM_BK;
    load_stage(0, 0);
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (nk > 1) load_stage(1, 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    LmCellIn cin[2][2];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int half = 0; half < 2; ++half)
            cin[m][half] = lm_cell_load(xproj, c_state, n0 + wm * 32 + m * 16 + gid + half * 8, N, T, t, H, dir, j);
    // ldmatrix lane addressing.  A (16 x 16, row-major): matrices (r0-7,k0-7), (r8-15,k0-7), (r0-7,k8-15), (r8-15,k8-15)
    const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_k = (lane >> 4) * 8;
    // B: per gate one x4 = hi(k0-7), hi(k8-15), lo(k0-7), lo(k8-15) of the warp's 8 units
    const int b_row = lane & 7, b_plane = lane >> 4, b_k = ((lane >> 3) & 1) * 8;
    for (int kc = 0; kc < nk; ++kc) {
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();
        if (kc + 2 < nk) load_stage(kc + 2, (kc + 2) % LM_STAGES);
        asm volatile("cp.async.commit_group;" ::: "memory");
        const uint32_t sa = sbase + (kc % LM_STAGES) * C::STAGE;
        const uint32_t sb = sa + 2 * C::A_PLANE;
#pragma unroll
        for (int kk = 0; kk < LM_BK / 16; ++kk) {
            uint32_t ah[2][4], al[2][4];
#pragma unroll
            for (int m = 0; m < 2; ++m) {
                const uint32_t addr = sa + (wm * 32 + m * 16 + a_row) * LM_ROWB + (kk * 16 + a_k) * 2;
                lm_ldsm4(addr, ah[m]);
                lm_ldsm4(addr + C::A_PLANE, al[m]);
            }
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                uint32_t b[4];      // hi b0, hi b1, lo b0, lo b1
                lm_ldsm4(sb + b_plane * C::B_PLANE + (g * C::JB + wn * 8 + b_row) * LM_ROWB + (kk * 16 + b_k) * 2, b);
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    lm_mma(acc[m][g], al[m], b[0], b[1]);
                    lm_mma(acc[m][g], ah[m], b[2], b[3]);
                    lm_mma(acc[m][g], ah[m], b[0], b[1]);
                }
            }
        }
    }
    // fused cell update (gate order i, f, g, o as torch.nn.LSTM): thread holds rows gid, gid+8 and units 2q, 2q+1 of each m-tile
    __nv_bfloat16* hn = h_next + (long)dir * 2 * plane_h;
#pragma unroll
    for (int m = 0; m < 2; ++m) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const long n = n0 + wm * 32 + m * 16 + gid + half * 8;
            if (n >= N) continue;
            lm_cell_store(cin[m][half], acc[m][0][2 * half], acc[m][0][2 * half + 1], acc[m][1][2 * half], acc[m][1][2 * half + 1],
                          acc[m][2][2 * half], acc[m][2][2 * half + 1], acc[m][3][2 * half], acc[m][3][2 * half + 1], c_state, out, hn,
                          plane_h, n, N, T, t, H, dir, j);
        }
    }
}

template <int WM, int WN>
int lstm_step_mma_launch(const float* xp, const __nv_bfloat16* whh, const __nv_bfloat16* hp, __nv_bfloat16* hn, float* cs, float* y,
                         long N, int T, int H, int s, cudaStream_t st) {
    using C = LmCfg<WM, WN>;
    static CssPerDevice attr;
    if (attr.first()) {
        cudaError_t e = cudaFuncSetAttribute(lstm_step_mma_kernel<WM, WN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        CSS_REQUIRE(e == cudaSuccess, "lstm_step_mma: cannot reserve %d B of shared memory: %s", C::SMEM, cudaGetErrorString(e));
    }
    dim3 grid((unsigned)((N + C::BM - 1) / C::BM), H / C::JB, 2);
    lstm_step_mma_kernel<WM, WN><<<grid, C::THREADS, C::SMEM, st>>>(xp, whh, hp, hn, cs, y, N, T, H, s);
    return 0;
}

// One 64-column group of the resident kernels: 4 k16 steps x (4 gate tiles x 3 product terms).  The fragments of step kk + 1
// are requested (6 ldmatrix.x4) before the 12 HMMAs of step kk are issued -- with one warp per scheduler the loop was
// otherwise one shared-memory round trip per gate tile (ncu: 35 cycles per HMMA, short-scoreboard + wait stalls) -- and the
// HMMAs go term by term over the four gate accumulators, so consecutive ones are independent.  Every accumulator still sees
// lo.hi, hi.lo, hi.hi in that order: the sums are the bits the ring kernel produces.
__device__ __forceinline__ void lm_group_mma(float (&acc)[4][4], int c_first, uint32_t a_base, uint32_t a_plane, int a_c, int a_row,
                                             uint32_t b_base, int pitch, int b_c, int b_row) {
    uint32_t fa[2][8], fb[2][16];
    auto fetch = [&](int buf, int c0) {
        const uint32_t aaddr = a_base + (((c0 + a_c) ^ (a_row & 7)) << 4);
        lm_ldsm4(aaddr, &fa[buf][0]);                        // hi plane of h
        lm_ldsm4(aaddr + a_plane, &fa[buf][4]);              // lo plane
        const uint32_t boff = ((c0 + b_c) ^ b_row) << 4;
#pragma unroll
        for (int gt = 0; gt < 4; ++gt) lm_ldsm4(b_base + gt * 8 * pitch + boff, &fb[buf][4 * gt]);   // hi b0, hi b1, lo b0, lo b1
    };
    fetch(0, c_first);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
        const int cur = kk & 1;
        if (kk < 3) fetch(cur ^ 1, c_first + (kk + 1) * 2);
#pragma unroll
        for (int gt = 0; gt < 4; ++gt) lm_mma(acc[gt], &fa[cur][4], fb[cur][4 * gt], fb[cur][4 * gt + 1]);
#pragma unroll
        for (int gt = 0; gt < 4; ++gt) lm_mma(acc[gt], &fa[cur][0], fb[cur][4 * gt + 2], fb[cur][4 * gt + 3]);
#pragma unroll
        for (int gt = 0; gt < 4; ++gt) lm_mma(acc[gt], &fa[cur][0], fb[cur][4 * gt], fb[cur][4 * gt + 1]);
    }
}

// Few sequences (one 60 s recording is 72 chunks; up to 160): every CTA takes ALL K of its operands into shared memory -- up to
// 80 sequences (five m16 warp tiles) x H of h_{t-1} and 8 hidden units x 4 gates x H of W_hh, hi and lo planes, 224 KB at
// H = 512 -- as H/64 cp.async groups that the mma loop waits on one by one (no ring, no reuse barrier).  Rows are H bf16
// wide without padding; 16-byte chunk c of row r sits at chunk c ^ (r & 7) (conflict-free ldmatrix).  grid.y = H/8 CTAs
// per direction: 128 CTAs at H = 512, one wave.
//
// The kernel runs steps [s_begin, s_end) of a layer.  One step per launch (N <= 160, two row groups) or -- N <= 80 -- the WHOLE
// recurrence in one cooperative launch: W_hh of the CTA (64 KB at H = 512) is copied into shared memory once and stays, the cell
// state of a thread's (sequence, unit) pairs stays in registers, per step only the planes of h_{t-1} come in (cp.async.cg from
// L2) and the planes of h_t go out, the two h buffers alternating.  The CTAs of one direction meet at a counter in global memory
// after every step (release: bar.sync + fence + atomic add by thread 0; acquire: relaxed spin + fence + bar.sync); the two
// directions never wait for each other.  The gate pre-activations of the next step are requested before the spin.  All CTAs
// must be co-resident: cooperative launch, grid <= one CTA per SM (checked by the host, which otherwise keeps the per-step
// launches).  A spin that sees no progress for ~1 s (a lost co-residency guarantee: preemption, a debugger) traps: the launch
// fails with a CUDA error at the next synchronisation instead of hanging the GPU or returning masks computed from stale h.
//
// KS = 2: two sets of MW warps, set k owns the 64-column groups g = k (mod 2) -- it copies them, waits for them behind its own
// named barrier and multiplies them -- so every scheduler carries 2-3 warps instead of 1-2 (five warps on four schedulers left
// one of them with twice the HMMAs).  The partial sums of set 1 cross to set 0 through the (then idle) h tile.
__device__ __forceinline__ void lm_x_load(LmCellIn& in, const float* __restrict__ xproj, long n, long N, int T, int t, int H, int dir, int j) {
    if (n >= N) n = N - 1;
    const float* xp = xproj + (n * T + t) * (long)(8 * H) + (long)dir * 4 * H + j;
    in.xi = __ldg(reinterpret_cast<const float2*>(xp));
    in.xf = __ldg(reinterpret_cast<const float2*>(xp + H));
    in.xg = __ldg(reinterpret_cast<const float2*>(xp + 2 * H));
    in.xo = __ldg(reinterpret_cast<const float2*>(xp + 3 * H));
}
__device__ __forceinline__ void lm_wait_groups(int n) {      // at most n cp.async groups of this thread still in flight
    switch (n) {
        case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
        case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
        case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
        case 3: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
        case 4: asm volatile("cp.async.wait_group 4;" ::: "memory"); break;
        case 5: asm volatile("cp.async.wait_group 5;" ::: "memory"); break;
        case 6: asm volatile("cp.async.wait_group 6;" ::: "memory"); break;
        default: asm volatile("cp.async.wait_group 7;" ::: "memory"); break;
    }
}

template <int MW, int KS>
__global__ void __launch_bounds__(32 * MW * KS)
lstm_res_kernel(const float* __restrict__ xproj, const __nv_bfloat16* __restrict__ whh, __nv_bfloat16* __restrict__ hbuf0,
                __nv_bfloat16* __restrict__ hbuf1, float* __restrict__ c_state, float* __restrict__ out, unsigned* __restrict__ sync,
                long N, int T, int H, int s_begin, int s_end) {
    constexpr int RM = 16 * MW, SET = 32 * MW;
    extern __shared__ __align__(128) uint8_t lm_smem[];
    const int dir = blockIdx.z;
    const long n0 = (long)blockIdx.x * RM;
    const int j0 = blockIdx.y * 8;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ks = warp / MW, mt = warp - ks * MW, tl = tid - ks * SET;    // k set, m tile, thread index inside the set
    const long plane_h = N * (long)H, plane_w = 4L * H * H;
    const __nv_bfloat16* wp = whh + (long)dir * 2 * plane_w;
    const int pitch = H * 2;
    const uint32_t sA = (uint32_t)__cvta_generic_to_shared(lm_smem);
    const uint32_t a_plane = RM * pitch, b_plane_b = 32 * pitch;
    const uint32_t sB = sA + 2 * a_plane;
    const int ngl = (H >> 6) / KS;                           // 64-column groups of this set: g = i * KS + ks
    // every thread copies the same (row, chunk-of-8) slots in each group: offsets are worked out once (the swizzle only touches
    // the low three chunk bits, which a group offset leaves alone)
    constexpr int A_IT = 2 * RM * 8 / SET, B_IT = (2 * 32 * 8 + SET - 1) / SET;
    long aoff[A_IT];                                         // element offset inside one h buffer
    uint32_t adst[A_IT];
    const int c7 = tl & 7;
#pragma unroll
    for (int it = 0; it < A_IT; ++it) {
        const int rs = (tl >> 3) + it * (SET / 8), p = rs / RM, r = rs - p * RM;
        long n = n0 + r;
        if (n >= N) n = N - 1;                               // rows past N are computed and dropped
        aoff[it] = (long)dir * 2 * plane_h + p * plane_h + n * H + c7 * 8;
        adst[it] = sA + p * a_plane + r * pitch + ((c7 ^ (r & 7)) << 4);
    }
    const bool live = n0 + mt * 16 < N;                      // warp-uniform: m-tiles entirely past N skip the products
    const bool owner = live && ks == 0;                      // the warp that holds the cell state and writes h_t
    const int gid = lane >> 2, q = lane & 3;
    const int j = j0 + 2 * q;
    const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_c = lane >> 4;
    const int b_row = lane & 7, b_pl = lane >> 4, b_c = (lane >> 3) & 1;
    const uint32_t a_base = sA + (mt * 16 + a_row) * pitch;
    const uint32_t b_base = sB + b_pl * b_plane_b + b_row * pitch;
    const unsigned ncta = gridDim.x * gridDim.y;             // CTAs of one direction
    unsigned* ctr = sync + dir;
    bool dead = false;
    LmCellIn cin[2];
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        long n = n0 + mt * 16 + gid + half * 8;
        if (n >= N) n = N - 1;
        cin[half].c = make_float2(0.f, 0.f);
        if (owner) {
            if (s_begin > 0) cin[half].c = *reinterpret_cast<const float2*>(c_state + ((long)dir * N + n) * H + j);
            lm_x_load(cin[half], xproj, n, N, T, dir ? T - 1 - s_begin : s_begin, H, dir, j);
        }
    }
    for (int s = s_begin; s < s_end; ++s) {
        const int t = dir ? (T - 1 - s) : s;
        const __nv_bfloat16* hp = (s & 1) ? hbuf1 : hbuf0;
        __nv_bfloat16* hn = ((s & 1) ? hbuf0 : hbuf1) + (long)dir * 2 * plane_h;
        for (int i = 0; i < ngl; ++i) {
            const int g = i * KS + ks;
#pragma unroll
            for (int it = 0; it < A_IT; ++it) lm_cp16(adst[it] + g * 128, hp + aoff[it] + g * 64);
            if (s == s_begin) {
#pragma unroll
                for (int it = 0; it < B_IT; ++it) {
                    const int rs = (tl >> 3) + it * (SET / 8);
                    if (rs < 64) {
                        const int p = rs >> 5, r = rs & 31;  // r = gate * 8 + unit
                        lm_cp16(sB + p * b_plane_b + r * pitch + ((c7 ^ (r & 7)) << 4) + g * 128,
                                wp + p * plane_w + ((long)(r >> 3) * H + j0 + (r & 7)) * H + c7 * 8 + g * 64);
                    }
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        float acc[4][4];
#pragma unroll
        for (int g = 0; g < 4; ++g)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[g][i] = 0.f;
        for (int i = 0; i < ngl; ++i) {
            lm_wait_groups(ngl - 1 - i);
            if (KS == 1) __syncthreads();
            else asm volatile("bar.sync %0, %1;" ::"r"(1 + ks), "n"(SET) : "memory");
            if (live) lm_group_mma(acc, (i * KS + ks) * 8, a_base, a_plane, a_c, a_row, b_base, pitch, b_c, b_row);
        }
        if (KS > 1) {                                        // partial sums of set 1 -> set 0, through the h tile
            float4* part = reinterpret_cast<float4*>(lm_smem);
            __syncthreads();                                 // both sets are done reading the h tile
            if (live && ks == 1) {
#pragma unroll
                for (int g = 0; g < 4; ++g) part[g * SET + tl] = make_float4(acc[g][0], acc[g][1], acc[g][2], acc[g][3]);
            }
            __syncthreads();
            if (owner) {
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const float4 v = part[g * SET + tl];
                    acc[g][0] += v.x; acc[g][1] += v.y; acc[g][2] += v.z; acc[g][3] += v.w;
                }
            }
        }
        if (owner) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const long n = n0 + mt * 16 + gid + half * 8;
                if (n >= N) continue;
                const LmCellIn& in = cin[half];
                const float ai0 = acc[0][2 * half], ai1 = acc[0][2 * half + 1], af0 = acc[1][2 * half], af1 = acc[1][2 * half + 1];
                const float ag0 = acc[2][2 * half], ag1 = acc[2][2 * half + 1], ao0 = acc[3][2 * half], ao1 = acc[3][2 * half + 1];
                float2 c, h;                                  // same expressions as lm_cell_store
                c.x = lm_sigmoid(af0 + in.xf.x) * in.c.x + lm_sigmoid(ai0 + in.xi.x) * lm_tanh(ag0 + in.xg.x);
                c.y = lm_sigmoid(af1 + in.xf.y) * in.c.y + lm_sigmoid(ai1 + in.xi.y) * lm_tanh(ag1 + in.xg.y);
                h.x = lm_sigmoid(ao0 + in.xo.x) * lm_tanh(c.x);
                h.y = lm_sigmoid(ao1 + in.xo.y) * lm_tanh(c.y);
                cin[half].c = c;
                *reinterpret_cast<float2*>(out + (n * T + t) * (long)(2 * H) + (long)dir * H + j) = h;
                const __nv_bfloat162 hi = __floats2bfloat162_rn(h.x, h.y);
                const __nv_bfloat162 lo = __floats2bfloat162_rn(h.x - __bfloat162float(hi.x), h.y - __bfloat162float(hi.y));
                *reinterpret_cast<__nv_bfloat162*>(hn + n * H + j) = hi;
                *reinterpret_cast<__nv_bfloat162*>(hn + plane_h + n * H + j) = lo;
            }
        }
        if (s + 1 == s_end) break;
        if (owner) {
#pragma unroll
            for (int half = 0; half < 2; ++half)
                lm_x_load(cin[half], xproj, n0 + mt * 16 + gid + half * 8, N, T, dir ? (T - 2 - s) : (s + 1), H, dir, j);
        }
        __syncthreads();                                     // this CTA's h_t stores are issued, its reads of the h tile are done
        if (tid == 0) {
            __threadfence();
            atomicAdd(ctr, 1u);
            if (!dead) {
                const unsigned target = (unsigned)(s - s_begin + 1) * ncta;
                const long long t0 = clock64();
                unsigned v;
                for (;;) {
                    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");   // the fence below orders the h loads
                    if (v >= target) break;
                    if (clock64() - t0 > 2000000000LL) { atomicExch(sync + 2, 1u); __trap(); }   // fail loudly (below)
                }
            }
            __threadfence();
        }
        __syncthreads();
    }
    if (owner && s_end < T) {                                // the next launch continues from here
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const long n = n0 + mt * 16 + gid + half * 8;
            if (n < N) *reinterpret_cast<float2*>(c_state + ((long)dir * N + n) * H + j) = cin[half].c;
        }
    }
}

// Whole recurrence, two warp groups.  The sequences of a batch are independent recurrences, so nothing ties the m tiles of a
// CTA to each other: group 0 (m tiles 0-1, rows 0-31) and group 1 (m tiles 2-4, rows 32-79) each have their rows of the h
// tile, their two k sets, their named barriers and their own step counter per direction, and drift apart -- while one group
// waits for its counter the other one multiplies, and the HMMAs of a group meet half the contention.  W_hh is copied once by
// all threads before the groups separate.  Same arithmetic per (sequence, unit) as lstm_res_kernel<5, 2>: equal bits.
__device__ __forceinline__ void lm_bar(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }

__global__ void __launch_bounds__(320)
lstm_res2_kernel(const float* __restrict__ xproj, const __nv_bfloat16* __restrict__ whh, __nv_bfloat16* __restrict__ hbuf0,
                 __nv_bfloat16* __restrict__ hbuf1, float* __restrict__ out, unsigned* __restrict__ sync, long N, int T, int H) {
    constexpr int RM = 80;
    extern __shared__ __align__(128) uint8_t lm_smem[];
    const int dir = blockIdx.z;
    const long n0 = (long)blockIdx.x * RM;
    const int j0 = blockIdx.y * 8;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int grp = warp >= 4;                               // warps 0-3: m tiles 0, 1; warps 4-9: m tiles 2, 3, 4
    const int n_mt = grp ? 3 : 2, gw0 = grp ? 4 : 0, row0 = grp ? 32 : 0, rows = n_mt * 16;
    const int wg = warp - gw0, ks = wg / n_mt, mt = (grp ? 2 : 0) + wg - ks * n_mt;
    const int SETN = 32 * n_mt, GRPN = 2 * SETN;
    const int tl = tid - (gw0 + ks * n_mt) * 32;             // thread index inside the k set
    const int bar_grp = 1 + grp, bar_set = 3 + 2 * grp + ks;
    const long plane_h = N * (long)H, plane_w = 4L * H * H;
    const __nv_bfloat16* wp = whh + (long)dir * 2 * plane_w;
    const int pitch = H * 2;
    const uint32_t sA = (uint32_t)__cvta_generic_to_shared(lm_smem);
    const uint32_t a_plane = RM * pitch, b_plane_b = 32 * pitch;
    const uint32_t sB = sA + 2 * a_plane;
    const int ngl = (H >> 6) / 2;                            // 64-column groups of this set: g = 2 i + ks
    // W_hh slice of the CTA: 64 row-planes x H/8 chunks, once, by everybody
    const int cpr = H >> 3;
    for (int idx = tid; idx < 64 * cpr; idx += 320) {
        const int rs = idx / cpr, c = idx - rs * cpr, p = rs >> 5, r = rs & 31;        // r = gate * 8 + unit
        lm_cp16(sB + p * b_plane_b + r * pitch + (((c & 7) ^ (r & 7)) << 4) + (c >> 3) * 128,
                wp + p * plane_w + ((long)(r >> 3) * H + j0 + (r & 7)) * H + c * 8);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    long aoff[8];
    uint32_t adst[8];
    const int c7 = tl & 7;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const int rs = (tl + it * SETN) >> 3, p = rs / rows, r = row0 + rs - p * rows;
        long n = n0 + r;
        if (n >= N) n = N - 1;                               // rows past N are computed and dropped
        aoff[it] = (long)dir * 2 * plane_h + p * plane_h + n * H + c7 * 8;
        adst[it] = sA + p * a_plane + r * pitch + ((c7 ^ (r & 7)) << 4);
    }
    const bool live = n0 + mt * 16 < N;
    const bool owner = live && ks == 0;
    const int gid = lane >> 2, q = lane & 3;
    const int j = j0 + 2 * q;
    const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_c = lane >> 4;
    const int b_row = lane & 7, b_pl = lane >> 4, b_c = (lane >> 3) & 1;
    const uint32_t a_base = sA + (mt * 16 + a_row) * pitch;
    const uint32_t b_base = sB + b_pl * b_plane_b + b_row * pitch;
    const unsigned ncta = gridDim.x * gridDim.y;
    unsigned* ctr = sync + dir * 2 + grp;
    float4* part = reinterpret_cast<float4*>(lm_smem + row0 * pitch);      // inside this group's rows of the h tile
    bool dead = false;
    LmCellIn cin[2];
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        cin[half].c = make_float2(0.f, 0.f);
        if (owner) lm_x_load(cin[half], xproj, n0 + mt * 16 + gid + half * 8, N, T, dir ? T - 1 : 0, H, dir, j);
    }
    for (int s = 0; s < T; ++s) {
        const int t = dir ? (T - 1 - s) : s;
        const __nv_bfloat16* hp = (s & 1) ? hbuf1 : hbuf0;
        __nv_bfloat16* hn = ((s & 1) ? hbuf0 : hbuf1) + (long)dir * 2 * plane_h;
        for (int i = 0; i < ngl; ++i) {
            const int g = i * 2 + ks;
#pragma unroll
            for (int it = 0; it < 8; ++it) lm_cp16(adst[it] + g * 128, hp + aoff[it] + g * 64);
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        float acc[4][4];
#pragma unroll
        for (int g = 0; g < 4; ++g)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[g][i] = 0.f;
        for (int i = 0; i < ngl; ++i) {
            lm_wait_groups(ngl - 1 - i);
            lm_bar(bar_set, SETN);
            if (live) lm_group_mma(acc, (i * 2 + ks) * 8, a_base, a_plane, a_c, a_row, b_base, pitch, b_c, b_row);
        }
        lm_bar(bar_grp, GRPN);                               // both sets are done reading this group's rows of the h tile
        if (live && ks == 1) {
#pragma unroll
            for (int g = 0; g < 4; ++g) part[g * SETN + tl] = make_float4(acc[g][0], acc[g][1], acc[g][2], acc[g][3]);
        }
        lm_bar(bar_grp, GRPN);
        if (owner) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const float4 v = part[g * SETN + tl];
                acc[g][0] += v.x; acc[g][1] += v.y; acc[g][2] += v.z; acc[g][3] += v.w;
            }
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const long n = n0 + mt * 16 + gid + half * 8;
                if (n >= N) continue;
                const LmCellIn& in = cin[half];
                const float ai0 = acc[0][2 * half], ai1 = acc[0][2 * half + 1], af0 = acc[1][2 * half], af1 = acc[1][2 * half + 1];
                const float ag0 = acc[2][2 * half], ag1 = acc[2][2 * half + 1], ao0 = acc[3][2 * half], ao1 = acc[3][2 * half + 1];
                float2 c, h;                                  // same expressions as lm_cell_store
                c.x = lm_sigmoid(af0 + in.xf.x) * in.c.x + lm_sigmoid(ai0 + in.xi.x) * lm_tanh(ag0 + in.xg.x);
                c.y = lm_sigmoid(af1 + in.xf.y) * in.c.y + lm_sigmoid(ai1 + in.xi.y) * lm_tanh(ag1 + in.xg.y);
                h.x = lm_sigmoid(ao0 + in.xo.x) * lm_tanh(c.x);
                h.y = lm_sigmoid(ao1 + in.xo.y) * lm_tanh(c.y);
                cin[half].c = c;
                *reinterpret_cast<float2*>(out + (n * T + t) * (long)(2 * H) + (long)dir * H + j) = h;
                const __nv_bfloat162 hi = __floats2bfloat162_rn(h.x, h.y);
                const __nv_bfloat162 lo = __floats2bfloat162_rn(h.x - __bfloat162float(hi.x), h.y - __bfloat162float(hi.y));
                *reinterpret_cast<__nv_bfloat162*>(hn + n * H + j) = hi;
                *reinterpret_cast<__nv_bfloat162*>(hn + plane_h + n * H + j) = lo;
            }
        }
        if (s + 1 == T) break;
        if (owner) {
#pragma unroll
            for (int half = 0; half < 2; ++half)
                lm_x_load(cin[half], xproj, n0 + mt * 16 + gid + half * 8, N, T, dir ? (T - 2 - s) : (s + 1), H, dir, j);
        }
        lm_bar(bar_grp, GRPN);                               // the group's h_t stores are issued, its reads of the partial sums are done
        if (wg == 0 && lane == 0) {
            __threadfence();
            atomicAdd(ctr, 1u);
            if (!dead) {
                const unsigned target = (unsigned)(s + 1) * ncta;
                const long long t0 = clock64();
                unsigned v;
                for (;;) {
                    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
                    if (v >= target) break;
                    if (clock64() - t0 > 2000000000LL) { atomicExch(sync + 8, 1u); __trap(); }
                }
            }
            __threadfence();
        }
        lm_bar(bar_grp, GRPN);
    }
}

int lstm_res2_launch(const float* xp, const __nv_bfloat16* whh, __nv_bfloat16* h0, __nv_bfloat16* h1, float* y, unsigned* sync, long N, int T,
                     int H, cudaStream_t st) {
    int smem = 2 * (80 + 32) * H * 2;
    static int attr_smem_dev[CSS_MAX_DEVICES] = {};
    int& attr_smem = attr_smem_dev[css_device()];
    if (smem > attr_smem) {
        cudaError_t e = cudaFuncSetAttribute(lstm_res2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        CSS_REQUIRE(e == cudaSuccess, "lstm_res2: cannot reserve %d B of shared memory: %s", smem, cudaGetErrorString(e));
        attr_smem = smem;
    }
    dim3 grid(1, H / 8, 2), block(320);
    void* args[] = {(void*)&xp, (void*)&whh, (void*)&h0, (void*)&h1, (void*)&y, (void*)&sync, (void*)&N, (void*)&T, (void*)&H};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)lstm_res2_kernel, grid, block, args, (size_t)smem, st);
    CSS_REQUIRE(e == cudaSuccess, "lstm_res2: cooperative launch failed: %s", cudaGetErrorString(e));
    return 0;
}

constexpr int LM_RES_MW = 5;
static int lstm_seq_res_ok(long N, int H) {                  // 1: the whole grid fits the device at once
    static int sms_dev[CSS_MAX_DEVICES] = {}, coop_dev[CSS_MAX_DEVICES] = {};
    const int dev = css_device();
    int &sms = sms_dev[dev], &coop = coop_dev[dev];
    if (!sms) {
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    return coop && H % 64 == 0 && H <= 512 && N <= 16 * LM_RES_MW && 2 * (H / 8) <= sms;
}

// steps [s_begin, s_end) of one layer; more than one step per launch needs lstm_seq_res_ok (cooperative launch, zeroed `sync`)
template <int KS>
int lstm_res_launch_ks(const float* xp, const __nv_bfloat16* whh, __nv_bfloat16* h0, __nv_bfloat16* h1, float* cs, float* y, unsigned* sync,
                       long N, int T, int H, int s_begin, int s_end, cudaStream_t st) {
    int smem = 2 * (16 * LM_RES_MW + 32) * H * 2;
    static int attr_smem_dev[CSS_MAX_DEVICES] = {};
    int& attr_smem = attr_smem_dev[css_device()];
    if (smem > attr_smem) {
        cudaError_t e = cudaFuncSetAttribute(lstm_res_kernel<LM_RES_MW, KS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        CSS_REQUIRE(e == cudaSuccess, "lstm_res: cannot reserve %d B of shared memory: %s", smem, cudaGetErrorString(e));
        attr_smem = smem;
    }
    dim3 grid((unsigned)((N + 16 * LM_RES_MW - 1) / (16 * LM_RES_MW)), H / 8, 2), block(32 * LM_RES_MW * KS);
    if (s_end - s_begin == 1) {
        lstm_res_kernel<LM_RES_MW, KS><<<grid, block, smem, st>>>(xp, whh, h0, h1, cs, y, sync, N, T, H, s_begin, s_end);
        return 0;
    }
    void* args[] = {(void*)&xp, (void*)&whh, (void*)&h0, (void*)&h1, (void*)&cs, (void*)&y, (void*)&sync, (void*)&N, (void*)&T, (void*)&H,
                    (void*)&s_begin, (void*)&s_end};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)lstm_res_kernel<LM_RES_MW, KS>, grid, block, args, (size_t)smem, st);
    CSS_REQUIRE(e == cudaSuccess, "lstm_res: cooperative launch failed: %s", cudaGetErrorString(e));
    return 0;
}
int lstm_res_launch(const float* xp, const __nv_bfloat16* whh, __nv_bfloat16* h0, __nv_bfloat16* h1, float* cs, float* y, unsigned* sync,
                    long N, int T, int H, int s_begin, int s_end, cudaStream_t st) {
    static const int ks_env = getenv("CSS_B200_LSTM_KS") ? atoi(getenv("CSS_B200_LSTM_KS")) : 2;      // tuning: 1 = one warp set
    if (ks_env == 2 && (H >> 6) % 2 == 0) return lstm_res_launch_ks<2>(xp, whh, h0, h1, cs, y, sync, N, T, H, s_begin, s_end, st);
    return lstm_res_launch_ks<1>(xp, whh, h0, h1, cs, y, sync, N, T, H, s_begin, s_end, st);
}
}  // namespace

struct BlstmModel {
    css_blstm_desc d;
    std::vector<css_blstm_layer> lay;
    void* whh_planes = nullptr;      // per layer: [dir][plane][4H][H] bf16 (built on the first tensor-core forward)
    PlaneMat embed_w{}, head_w{};    // Linear weights as bf16 hi/lo planes for the tcgen05 GEMM (same allocation)
    std::vector<PlaneMat> w_ih;
};
extern "C" int css_get_gemm_mode(void);

static size_t whh_layer_elems(const css_blstm_desc& d) { const size_t H = d.hdim / 2; return 2 * 2 * 4 * H * H; }

static inline int embed_kp(const css_blstm_desc& d) { return (d.ldin + 63) / 64 * 64; }    // K of the embedding GEMM (zero padded)

static int build_whh_planes(BlstmModel* m, cudaStream_t st) {
    if (m->whh_planes) return 0;
    const css_blstm_desc& c = m->d;
    const size_t per = whh_layer_elems(c);
    const int H = c.hdim / 2;
    const size_t h = c.hdim, nh = (size_t)c.num_bins * c.num_out, kpe = embed_kp(c);
    auto bytes = [](size_t rows, size_t k) { return (2 * rows * k * 2 + 255) & ~(size_t)255; };
    const size_t whh_bytes = (per * c.num_layers * 2 + 255) & ~(size_t)255;
    const size_t total = whh_bytes + bytes(h, kpe) + bytes(nh, h) + c.num_layers * bytes(4 * h, h);
    cudaError_t e = cudaMalloc(&m->whh_planes, total);
    CSS_REQUIRE(e == cudaSuccess, "blstm: cannot allocate %zu B for the bf16 weight planes: %s", total, cudaGetErrorString(e));
    int rc = 0;
    char* wp = static_cast<char*>(m->whh_planes) + whh_bytes;
    auto mk = [&](const float* w, size_t rows, size_t cols, size_t k) {
        PlaneMat pm = {wp, (long)(rows * k), (int)k, 2};
        wp += bytes(rows, k);
        if (!rc) rc = split_planes_launch(w, (long)rows, (int)cols, (int)cols, pm, st);
        return pm;
    };
    m->embed_w = mk(c.embed_w, h, c.ldin, kpe);
    m->head_w = mk(c.head_w, nh, h, h);
    m->w_ih.clear();
    for (int l = 0; l < c.num_layers; ++l) m->w_ih.push_back(mk(m->lay[l].w_ih, 4 * h, h, h));
    for (int l = 0; l < m->d.num_layers && !rc; ++l)
        for (int dir = 0; dir < 2 && !rc; ++dir) {
            __nv_bfloat16* base = static_cast<__nv_bfloat16*>(m->whh_planes) + per * l + (per / 2) * dir;
            PlaneMat pm = {base, 4L * H * H, H, 2};
            rc = split_planes_launch(dir ? m->lay[l].w_hh_b : m->lay[l].w_hh_f, 4L * H, H, H, pm, st);
        }
    if (rc) { cudaFree(m->whh_planes); m->whh_planes = nullptr; }
    return rc;
}

static int g_lstm_plan = -1;
extern "C" int css_get_lstm_plan(void) {
    if (g_lstm_plan < 0) g_lstm_plan = getenv("CSS_B200_LSTM_TILE") ? atoi(getenv("CSS_B200_LSTM_TILE")) : 0;
    return g_lstm_plan;
}
extern "C" int css_set_lstm_plan(int plan) {
    CSS_REQUIRE(plan == 0 || plan == 1 || plan == 3 || plan == 12 || plan == 22, "css_set_lstm_plan: unknown plan %d", plan);
    g_lstm_plan = plan;
    return 0;
}

extern "C" void* css_blstm_create(const css_blstm_desc* desc, const css_blstm_layer* layers) {
    if (!desc || !layers) { css_set_error("css_blstm_create: null descriptor"); return nullptr; }
    if (desc->hdim % 128 || desc->hdim > 1024 || (desc->hdim / 2) % LS_BJ || desc->ldin % 16 || desc->ldin < desc->idim) {
        css_set_error("css_blstm_create: unsupported hdim=%d / ldin=%d", desc->hdim, desc->ldin);
        return nullptr;
    }
    BlstmModel* m = new (std::nothrow) BlstmModel();
    if (!m) { css_set_error("css_blstm_create: out of host memory"); return nullptr; }
    m->d = *desc;
    m->lay.assign(layers, layers + desc->num_layers);
    return m;
}
extern "C" void css_blstm_destroy(void* model) {
    BlstmModel* m = static_cast<BlstmModel*>(model);
    if (m && m->whh_planes) cudaFree(m->whh_planes);
    delete m;
}

static inline size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }
static inline int bl_head_pitch(const css_blstm_desc& d) { return ((d.num_bins * d.num_out + 7) / 8) * 8; }

extern "C" size_t css_blstm_workspace_bytes(const void* model, long N, int T) {
    const BlstmModel* m = static_cast<const BlstmModel*>(model);
    const size_t M = (size_t)N * T, h = m->d.hdim;
    // x, y, xproj, h (two fp32 state buffers, reused as the two bf16 hi/lo plane sets of the tensor-core step), c, head
    // ..., and the bf16 hi/lo planes of the current GEMM input (M x max(h, padded idim))
    const size_t ka = h > (size_t)embed_kp(m->d) ? h : (size_t)embed_kp(m->d);
    return 2 * al256(M * h * 4) + al256(M * 4 * h * 4) + al256((size_t)N * h * 4) * 3 +
           al256(M * bl_head_pitch(m->d) * 4) + al256(M * ka * 4);
}

extern "C" int css_blstm_forward(void* model, float* feat, long N, int T, int apply_mvn, float clamp_max, float* masks,
                                 float* head, int ld_head, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    BlstmModel* m = static_cast<BlstmModel*>(model);
    CSS_REQUIRE(m && feat && workspace, "css_blstm_forward: null argument");
    CSS_REQUIRE(N > 0 && T > 1, "css_blstm_forward: empty input (N=%ld, T=%d)", N, T);
    CSS_REQUIRE((masks != nullptr) != (head != nullptr), "css_blstm_forward: pass exactly one of masks/head");
    CSS_REQUIRE(workspace_bytes >= css_blstm_workspace_bytes(model, N, T), "css_blstm_forward: workspace too small");
    const css_blstm_desc& c = m->d;
    const long M = N * T;
    const int h = c.hdim, H = h / 2, nh = c.num_bins * c.num_out;
    char* wp = static_cast<char*>(workspace);
    float* x = (float*)wp;   wp += al256((size_t)M * h * 4);
    float* y = (float*)wp;   wp += al256((size_t)M * h * 4);
    float* xp = (float*)wp;  wp += al256((size_t)M * 4 * h * 4);
    float* hs0 = (float*)wp; wp += al256((size_t)N * h * 4);
    float* hs1 = (float*)wp; wp += al256((size_t)N * h * 4);
    float* cs = (float*)wp;  wp += al256((size_t)N * h * 4);
    float* hd = (float*)wp;  wp += al256((size_t)M * bl_head_pitch(c) * 4);
    void* act = wp;
    int rc;
    const bool tc = css_get_gemm_mode() != 0;
    const int P = css_get_gemm_mode() == 2 ? 1 : 2;
    // GEMM: tensor-core modes split the fp32 input into planes and run the tcgen05 kernel, mode 0 the fp32 SIMT kernel
    auto linear = [&](const float* A, int lda, int K, const float* W, const PlaneMat& Wp, float* C, int ldc, int Nn, const GemmEpilogue& ep) -> int {
        if (!tc) return gemm_dispatch(A, lda, W, lda, C, ldc, M, Nn, K, ep, st);
        PlaneMat aP = {act, M * (long)Wp.ld, Wp.ld, 2};
        int r = split_planes_launch(A, M, K, lda, aP, st);
        if (r) return r;
        css_prof_begin(0, 2.0 * (double)M * Nn * K, st);
        r = gemm_tc_planes(aP, Wp, P, M, Nn, Wp.ld, ep, C, ldc, nullptr, st);
        css_prof_end(st);
        return r;
    };
    const int lm_tile = css_get_lstm_plan();   // 1 resident per step, 3 resident whole sequence, 12 / 22 = WM, WN of the ring kernel
    if (tc && (rc = build_whh_planes(m, st))) return rc;
    if (apply_mvn && (rc = css_mvn_time(feat, N, T, c.idim, c.ldin, st))) return rc;
    GemmEpilogue ee = {c.embed_b, 0, 1.f, nullptr, 0, 0.f};
    if ((rc = linear(feat, c.ldin, c.ldin, c.embed_w, m->embed_w, y, h, h, ee))) return rc;
    if ((rc = layernorm_launch(y, c.embed_ln_g, c.embed_ln_b, x, M, h, 1e-5f, 1, st))) return rc;
    for (int l = 0; l < c.num_layers; ++l) {
        const css_blstm_layer& L = m->lay[l];
        GemmEpilogue ep = {L.b_gates, 0, 1.f, nullptr, 0, 0.f};
        if ((rc = linear(x, h, h, L.w_ih, tc ? m->w_ih[l] : PlaneMat{}, xp, 4 * h, 4 * h, ep))) return rc;
        cudaMemsetAsync(hs0, 0, (size_t)N * h * 4, st);
        cudaMemsetAsync(cs, 0, (size_t)N * h * 4, st);
        dim3 grid((unsigned)((N + LS_BN - 1) / LS_BN), H / LS_BJ, 2);
        float *hp = hs0, *hn = hs1;
        // tensor-core step: [dir][plane][N][H] bf16 = N*h*4 bytes, the size of one fp32 state buffer
        const __nv_bfloat16* whh = tc ? static_cast<const __nv_bfloat16*>(m->whh_planes) + whh_layer_elems(c) * l : nullptr;
        // few sequences: the T steps of the layer in one cooperative launch (0 = default; 1 forces the per-step launches)
        if (tc && (lm_tile == 0 || lm_tile == 3) && lstm_seq_res_ok(N, H)) {
            css_prof_begin(4, (double)T * 2.0 * 2.0 * (double)N * H * 4 * H, st);
            cudaMemsetAsync(hd, 0, 64, st);                  // step counters (the head buffer is not in use yet)
            static const int two_groups = getenv("CSS_B200_LSTM_GROUPS") ? atoi(getenv("CSS_B200_LSTM_GROUPS")) == 2 : 1;   // B200, 72 x 512: 3.47 -> 3.29 ms per 3 layers
            if (two_groups && (H >> 6) % 2 == 0 && N > 32)
                rc = lstm_res2_launch(xp, whh, reinterpret_cast<__nv_bfloat16*>(hs0), reinterpret_cast<__nv_bfloat16*>(hs1), y,
                                      reinterpret_cast<unsigned*>(hd), N, T, H, st);
            else
                rc = lstm_res_launch(xp, whh, reinterpret_cast<__nv_bfloat16*>(hs0), reinterpret_cast<__nv_bfloat16*>(hs1), cs, y,
                                     reinterpret_cast<unsigned*>(hd), N, T, H, 0, T, st);
            css_prof_end(st);
            if (rc || (rc = css_check_launch("lstm_seq"))) return rc;
            if ((rc = layernorm_launch(y, L.ln_g, L.ln_b, x, M, h, 1e-5f, 0, st))) return rc;
            continue;
        }
        for (int s = 0; s < T; ++s) {
            css_prof_begin(4, 2.0 * 2.0 * (double)N * H * 4 * H, st);
            if (tc) {
                __nv_bfloat16 *bp = reinterpret_cast<__nv_bfloat16*>(hp), *bn = reinterpret_cast<__nv_bfloat16*>(hn);
                // measured on B200 (N = 747, H = 512): 64 x 64 CTAs 49 us per step, 128 x 128 CTAs 58 us (1.3 waves of one CTA per SM)
                switch (lm_tile ? lm_tile : (N <= 160 && H <= 512 ? 1 : (N <= 192 ? 12 : 22))) {
                    case 1: rc = lstm_res_launch(xp, whh, reinterpret_cast<__nv_bfloat16*>(hs0), reinterpret_cast<__nv_bfloat16*>(hs1), cs, y, nullptr, N, T, H, s, s + 1, st); break;
                    case 12: rc = lstm_step_mma_launch<1, 2>(xp, whh, bp, bn, cs, y, N, T, H, s, st); break;
                    default: rc = lstm_step_mma_launch<2, 2>(xp, whh, bp, bn, cs, y, N, T, H, s, st); break;
                }
                if (rc) { css_prof_end(st); return rc; }
            } else {
                lstm_step_kernel<<<grid, 256, 0, st>>>(xp, L.w_hh_f, L.w_hh_b, hp, hn, cs, y, N, T, H, s);
            }
            css_prof_end(st);
            if ((rc = css_check_launch("lstm_step"))) return rc;
            float* tmp = hp; hp = hn; hn = tmp;
        }
        if ((rc = layernorm_launch(y, L.ln_g, L.ln_b, x, M, h, 1e-5f, 0, st))) return rc;
    }
    if (head) {
        GemmEpilogue eh = {c.head_b, 1, 1.f, nullptr, 0, 0.f};
        return linear(x, h, h, c.head_w, m->head_w, head, ld_head, nh, eh);
    }
    const int ldh = bl_head_pitch(c);
    GemmEpilogue eh = {c.head_b, 1, 1.f, nullptr, 0, clamp_max};
    if ((rc = linear(x, h, h, c.head_w, m->head_w, hd, ldh, nh, eh))) return rc;
    return masks_layout_launch(hd, ldh, masks, N, T, c.num_bins, c.num_out, st);
}
