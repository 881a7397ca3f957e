// Warp-level kernels of the Conformer block: LayerNorm, relative-position MHSA, conv module,
// and the mask-head layout transform.  (css/models/conformer.py:188-303, css/executor/separator.py:100-104)
#include "common.cuh"
#include <cstdlib>
#include "conformer.h"
#include "gemm.h"
#include <cuda_bf16.h>

// split 4 fp32 values into bf16 planes: plane p receives the bf16 rounding of what planes < p left over
__device__ __forceinline__ void store_split4(float4 o, __nv_bfloat16* base, long plane_stride, int n_planes) {
    float v[4] = {o.x, o.y, o.z, o.w};
    for (int p = 0; p < n_planes; ++p) {
        unsigned short h[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const __nv_bfloat16 b = __float2bfloat16_rn(v[e]);
            v[e] -= __bfloat162float(b);
            h[e] = __bfloat16_as_ushort(b);
        }
        *reinterpret_cast<uint2*>(base + p * plane_stride) =
            make_uint2((unsigned)h[0] | ((unsigned)h[1] << 16), (unsigned)h[2] | ((unsigned)h[3] << 16));
    }
}

// ---------------------------------------------------------------------------------------------
// LayerNorm over d (multiple of 128, <= 1024): one warp per row.  Optional fused ReLU.
// ---------------------------------------------------------------------------------------------
template <int VPL>   // float4 vectors per lane: d = 128 * VPL
__global__ void __launch_bounds__(256)
layernorm_kernel(const float* x, const float* __restrict__ g, const float* __restrict__ b, float* y,
                 long M, float eps, int relu, __nv_bfloat16* planes, long plane_stride, int n_planes,
                 const float* __restrict__ g2, const float* __restrict__ b2) {
    const long row = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= M) return;
    const int d = 128 * VPL;
    const float4* xr = reinterpret_cast<const float4*>(x + row * d);
    float4 v[VPL];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        v[i] = xr[lane + 32 * i];
        s += v[i].x + v[i].y + v[i].z + v[i].w;
    }
    const float mean = warp_sum(s) / d;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        float a = v[i].x - mean, bb = v[i].y - mean, c = v[i].z - mean, dd = v[i].w - mean;
        q += a * a + bb * bb + c * c + dd * dd;
    }
    const float rstd = rsqrtf(warp_sum(q) / d + eps);
    float4* yr = reinterpret_cast<float4*>(y + row * d);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const float4 gg = reinterpret_cast<const float4*>(g)[lane + 32 * i];
        const float4 bb = reinterpret_cast<const float4*>(b)[lane + 32 * i];
        float4 o;
        o.x = (v[i].x - mean) * rstd * gg.x + bb.x;
        o.y = (v[i].y - mean) * rstd * gg.y + bb.y;
        o.z = (v[i].z - mean) * rstd * gg.z + bb.z;
        o.w = (v[i].w - mean) * rstd * gg.w + bb.w;
        if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
        if (y) yr[lane + 32 * i] = o;
        if (planes && !g2) store_split4(o, planes + row * d + 4 * (lane + 32 * i), plane_stride, n_planes);
        v[i] = o;
    }
    if (!g2) return;
    // second LayerNorm of the row just produced (a block's output norm followed by the next block's FFN-in norm): the
    // planes hold LN2(LN1(x)), y holds LN1(x) -- the same values two launches would give, without re-reading y
    s = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) s += v[i].x + v[i].y + v[i].z + v[i].w;
    const float mean2 = warp_sum(s) / d;
    q = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        float a = v[i].x - mean2, bb = v[i].y - mean2, c = v[i].z - mean2, dd = v[i].w - mean2;
        q += a * a + bb * bb + c * c + dd * dd;
    }
    const float rstd2 = rsqrtf(warp_sum(q) / d + eps);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const float4 gg = reinterpret_cast<const float4*>(g2)[lane + 32 * i];
        const float4 bb = reinterpret_cast<const float4*>(b2)[lane + 32 * i];
        float4 o;
        o.x = (v[i].x - mean2) * rstd2 * gg.x + bb.x;
        o.y = (v[i].y - mean2) * rstd2 * gg.y + bb.y;
        o.z = (v[i].z - mean2) * rstd2 * gg.z + bb.z;
        o.w = (v[i].w - mean2) * rstd2 * gg.w + bb.w;
        store_split4(o, planes + row * d + 4 * (lane + 32 * i), plane_stride, n_planes);
    }
}

int layernorm_launch(const float* x, const float* g, const float* b, float* y, long M, int d, float eps, int relu,
                     cudaStream_t st) {
    return layernorm_planes_launch(x, g, b, y, M, d, eps, relu, nullptr, 0, 0, st);
}

// y (fp32, optional) and/or bf16 split planes (pitch d) of LayerNorm(x)
int layernorm_planes_launch(const float* x, const float* g, const float* b, float* y, long M, int d, float eps, int relu,
                            void* planes_v, long plane_stride, int n_planes, cudaStream_t st) {
    return layernorm2_planes_launch(x, g, b, y, nullptr, nullptr, M, d, eps, relu, planes_v, plane_stride, n_planes, st);
}

// y = LN(x; g, b) (fp32, optional); planes = split planes of LN(y; g2, b2) when g2 is given, of y otherwise
int layernorm2_planes_launch(const float* x, const float* g, const float* b, float* y, const float* g2, const float* b2, long M,
                             int d, float eps, int relu, void* planes_v, long plane_stride, int n_planes, cudaStream_t st) {
    __nv_bfloat16* planes = static_cast<__nv_bfloat16*>(planes_v);
    CSS_REQUIRE(M > 0 && (y || planes), "layernorm: empty input or no output");
    CSS_REQUIRE(!g2 || (b2 && planes), "layernorm: the second norm needs its bias and a plane output");
    CSS_REQUIRE(d % 128 == 0 && d <= 1024, "layernorm: d=%d must be a multiple of 128 and <= 1024", d);
    const unsigned grid = (unsigned)((M + 7) / 8);
    css_prof_begin(3, (double)M * d * (4.0 + (y ? 4.0 : 0.0) + 2.0 * (planes ? n_planes : 0)), st);
    switch (d / 128) {
        case 1: layernorm_kernel<1><<<grid, 256, 0, st>>>(x, g, b, y, M, eps, relu, planes, plane_stride, n_planes, g2, b2); break;
        case 2: layernorm_kernel<2><<<grid, 256, 0, st>>>(x, g, b, y, M, eps, relu, planes, plane_stride, n_planes, g2, b2); break;
        case 4: layernorm_kernel<4><<<grid, 256, 0, st>>>(x, g, b, y, M, eps, relu, planes, plane_stride, n_planes, g2, b2); break;
        case 8: layernorm_kernel<8><<<grid, 256, 0, st>>>(x, g, b, y, M, eps, relu, planes, plane_stride, n_planes, g2, b2); break;
        default: CSS_REQUIRE(false, "layernorm: unsupported d=%d", d);
    }
    css_prof_end(st);
    return css_check_launch("layernorm");
}

extern "C" int css_layernorm(const float* x, const float* g, const float* b, float* y, long M, int d, float eps,
                             int relu, cudaStream_t st) {
    return layernorm_launch(x, g, b, y, M, d, eps, relu, st);
}

// ---------------------------------------------------------------------------------------------
// Relative-position multi-head self-attention core (css/models/conformer.py:229-250):
//   scores[i][j] = q_i . (k_j + pe[i - j + maxlen]) / sqrt(dk) ; softmax_j ; out_i = sum_j p_ij v_j
// One CTA per (sequence, head).  dk = 64.  K and the needed 2T-1 rows of the embedding table are
// kept transposed in shared memory so that lanes (over j) read conflict-free.
// qkv: (N*T, 3*H*64) rows = [q | k | v], out: (N*T, H*64).
// ---------------------------------------------------------------------------------------------
#define ATT_DK 64
#define ATT_WARPS 8

__global__ void __launch_bounds__(ATT_WARPS * 32)
relpos_attention_kernel(const float* __restrict__ qkv, const float* __restrict__ pe, int pe_rows, int use_pe,
                        float* __restrict__ out, int T, int H, int v_smem) {
    extern __shared__ float sm[];
    const int TP = (T + 31) & ~31;             // padded j extent
    const int RP = 2 * TP;                     // padded relative extent
    float* Kt = sm;                            // [64][TP]
    float* Vs = Kt + ATT_DK * TP;              // [T][64]
    float* Pt = Vs + (v_smem ? T * ATT_DK : 0);   // [64][RP]   Pt[d][r] = pe[clamp(r - (T-1)) + maxlen][d]
    float* Pb = Pt + ATT_DK * RP;              // [warps][TP] probabilities
    float* Qb = Pb + ATT_WARPS * TP;           // [warps][64]
    const int n = blockIdx.x / H, h = blockIdx.x % H;
    const int ld = 3 * H * ATT_DK;
    const float* base = qkv + (long)n * T * ld;
    const int maxlen = pe_rows / 2;
    for (int i = threadIdx.x; i < T * ATT_DK; i += blockDim.x) {
        const int j = i >> 6, d = i & 63;
        Kt[d * TP + j] = base[(long)j * ld + (H + h) * ATT_DK + d];
        if (v_smem) Vs[j * ATT_DK + d] = base[(long)j * ld + (2 * H + h) * ATT_DK + d];
    }
    for (int i = threadIdx.x; i < ATT_DK * (TP - T); i += blockDim.x) {
        const int d = i / (TP - T), j = T + i % (TP - T);
        Kt[d * TP + j] = 0.f;
    }
    for (int i = threadIdx.x; i < RP * ATT_DK; i += blockDim.x) {
        const int r = i >> 6, d = i & 63;
        float v = 0.f;
        if (use_pe && r <= 2 * (T - 1)) {
            int rel = r - (T - 1);
            rel = max(-maxlen, min(maxlen - 1, rel)) + maxlen;
            v = pe[(long)rel * ATT_DK + d];
        }
        Pt[d * RP + r] = v;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* pb = Pb + warp * TP;
    float* qb = Qb + warp * ATT_DK;
    const int NJ = TP / 32;
    // V rows: shared memory when it fits (T <= ~190), else straight from global (L1/L2 resident)
    const float* Vsrc = v_smem ? Vs : base + (2 * H + h) * ATT_DK;
    const int vld = v_smem ? ATT_DK : ld;
    for (int i = warp; i < T; i += ATT_WARPS) {
        const float* qrow = base + (long)i * ld + h * ATT_DK;
        qb[lane] = qrow[lane];
        qb[lane + 32] = qrow[lane + 32];
        __syncwarp();
        float s[8];
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) s[jj] = 0.f;
        for (int d = 0; d < ATT_DK; ++d) {
            const float qd = qb[d];
            const float* kr = Kt + d * TP + lane;
            const float* pr = Pt + d * RP + (i + T - 1) - lane;    // r = i - j + T - 1
#pragma unroll
            for (int jj = 0; jj < 8; ++jj)
                if (jj < NJ) {
                    const int r = (i + T - 1) - lane - 32 * jj;
                    const float pv = (r >= 0) ? pr[-32 * jj] : 0.f;
                    s[jj] = fmaf(qd, kr[32 * jj] + pv, s[jj]);
                }
        }
        float mx = -INFINITY;
#pragma unroll
        for (int jj = 0; jj < 8; ++jj)
            if (jj < NJ) {
                s[jj] = (lane + 32 * jj < T) ? s[jj] * 0.125f : -INFINITY;
                mx = fmaxf(mx, s[jj]);
            }
        mx = warp_max(mx);
        float sum = 0.f;
#pragma unroll
        for (int jj = 0; jj < 8; ++jj)
            if (jj < NJ) {
                s[jj] = (lane + 32 * jj < T) ? expf(s[jj] - mx) : 0.f;
                sum += s[jj];
            }
        const float inv = 1.f / warp_sum(sum);
#pragma unroll
        for (int jj = 0; jj < 8; ++jj)
            if (jj < NJ) pb[lane + 32 * jj] = s[jj] * inv;
        __syncwarp();
        float o0 = 0.f, o1 = 0.f;
        for (int j = 0; j < T; ++j) {
            const float p = pb[j];
            o0 = fmaf(p, Vsrc[(long)j * vld + lane], o0);
            o1 = fmaf(p, Vsrc[(long)j * vld + lane + 32], o1);
        }
        float* orow = out + ((long)n * T + i) * (H * ATT_DK) + h * ATT_DK;
        orow[lane] = o0;
        orow[lane + 32] = o1;
        __syncwarp();
    }
}

size_t attention_smem_bytes(int T, int v_smem) {
    const int TP = (T + 31) & ~31;
    return sizeof(float) * ((size_t)ATT_DK * TP + (v_smem ? (size_t)T * ATT_DK : 0) + (size_t)ATT_DK * 2 * TP +
                            ATT_WARPS * TP + ATT_WARPS * ATT_DK);
}

int attention_launch(const float* qkv, const float* pe, int pe_rows, float* out, long N, int T, int H, int dk,
                     cudaStream_t st) {
    CSS_REQUIRE(dk == ATT_DK, "attention: head dim %d unsupported (kernel is built for 64)", dk);
    CSS_REQUIRE(N > 0 && T > 0 && T <= 256, "attention: T=%d out of range (1..256)", T);
    const int v_smem = attention_smem_bytes(T, 1) <= 227 * 1024;
    const size_t smem = attention_smem_bytes(T, v_smem);
    CSS_REQUIRE(smem <= 227 * 1024, "attention: T=%d needs %zu B of shared memory", T, smem);
    static CssPerDevice attr_set;
    if (attr_set.first()) {
        cudaFuncSetAttribute(relpos_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    }
    css_prof_begin(1, 6.0 * (double)N * H * T * T * ATT_DK, st);   // QK^T + Q.PE^T + PV
    relpos_attention_kernel<<<(unsigned)(N * H), ATT_WARPS * 32, smem, st>>>(qkv, pe, pe_rows, pe != nullptr, out, T, H, v_smem);
    css_prof_end(st);
    return css_check_launch("relpos_attention");
}

extern "C" int css_relpos_attention(const float* qkv, const float* pe, int pe_rows, float* out, long N, int T, int H,
                                    int dk, cudaStream_t st) {
    return attention_launch(qkv, pe, pe_rows, out, N, T, H, dk, st);
}

// ---------------------------------------------------------------------------------------------
// Conv module (css/models/conformer.py:289-303), fully fused:
//   LN -> scalar GLU (a*sigmoid(g), both scalar affines of the LN output) -> depthwise conv over
//   time (k taps, zero padded) -> BatchNorm (folded running stats) -> ReLU -> scalar affine -> + x.
// One CTA per (sequence, time tile); the gated activations of the tile plus a 16-row halo on each
// side live in shared memory; thread c then slides the taps of channel c down the time axis.
// ---------------------------------------------------------------------------------------------
#define CM_D 256
#define CM_HALO 16
#define CM_TAPS (2 * CM_HALO + 1)
#define CM_THREADS 512
#define CM_MAX_ROWS 192          // (tile + 2*halo) rows of 1 KB each + the tap table must fit in shared memory
#define CM_WPITCH (CM_D + 1)     // tap table in shared memory: [tap][channel], pitch 257 floats
#define CM_SPLIT_ROWS 112        // two-CTAs-per-SM shape: 112 KB per CTA
#define CM_OUT 8                 // outputs per register pass of the depthwise filter

// One CTA per (sequence, time tile); a 150-frame chunk is one tile, so nothing is recomputed.
// Phase 1: 16 warps, one row each: LayerNorm + scalar GLU into shared memory (zero rows outside [0, T)).
// Phase 2: thread = (channel pair, quarter of the tile); 8 outputs per pass from 40 shared-memory reads (5 per output
// instead of 33), taps and accumulators in registers as channel pairs so that every filter step is one packed
// fp32 FMA (FFMA2) -- the kernel is bound by the FMA pipe, not by memory.
// Tiling (ConvPlan): the first n_full sequences get `full_tiles` tiles each; what is left -- less than one wave of CTAs,
// which would otherwise cost a whole extra wave -- is cut into `tail_tiles` short tiles spread over the idle SMs.
struct ConvPlan { int n_full, full_tiles, full_len, tail_tiles, tail_len, smem_rows; };

// Two shapes: THREADS = 512, one CTA per SM, a whole 150-frame chunk per tile and the taps transposed into shared memory by the
// kernel; THREADS = 256, TWO CTAs per SM on half-chunk tiles (75 + 32 halo rows, 107 KB) with the taps read from a table the
// caller has transposed once (ConvModuleParams::dw_wt) -- one CTA's LayerNorm phase (loads) then runs under the other's filter
// phase (FMA pipe), at the price of normalising the 32 halo rows twice.
template <int THREADS>
__global__ void __launch_bounds__(THREADS, THREADS == 512 ? 1 : 2)
conv_module_kernel(const float* __restrict__ x, float* __restrict__ out, int T, ConvPlan pl, ConvModuleParams p) {
    constexpr bool WSMEM = THREADS == 512;
    extern __shared__ float U[];   // [(tile + 2*CM_HALO)][CM_D], then (WSMEM) the tap table [CM_TAPS][CM_WPITCH]
    long n;
    int t0, tile;
    {
        const int bid = blockIdx.x, nf = pl.n_full * pl.full_tiles;
        if (bid < nf) {
            n = bid / pl.full_tiles;
            t0 = (bid - (int)n * pl.full_tiles) * pl.full_len;
            tile = min(pl.full_len, T - t0);
        } else {
            const int r = bid - nf, q = r / pl.tail_tiles;
            n = pl.n_full + q;
            t0 = (r - q * pl.tail_tiles) * pl.tail_len;
            tile = min(pl.tail_len, T - t0);
        }
    }
    const int rows = tile + 2 * CM_HALO;
    const float* xs = x + n * (long)T * CM_D;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // The depthwise taps (d, 33) go to shared memory transposed, by 4-byte async copies that run under phase 1: read in
    // place by thread = channel they were 66 loads with a 132-byte stride between lanes, a fifth of the kernel's stall samples.
    float* Wt = U + (size_t)pl.smem_rows * CM_D;
    if (WSMEM) {
        for (int i = threadIdx.x; i < CM_D * CM_TAPS; i += THREADS) {
            const int ch = i / CM_TAPS, k = i - ch * CM_TAPS;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(Wt + k * CM_WPITCH + ch);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(p.dw_w + i) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    // four rows per pass: their eight loads are in flight together (the phase is latency-bound otherwise)
    constexpr int NW = THREADS / 32, RB = 4;
    for (int rb = warp; rb < rows; rb += NW * RB) {
        float4 v[RB][2];
#pragma unroll
        for (int i = 0; i < RB; ++i) {
            const int r = rb + i * NW, t = t0 - CM_HALO + r;
            if (r < rows && t >= 0 && t < T) {
                const float4* xr = reinterpret_cast<const float4*>(xs + (long)t * CM_D);
                v[i][0] = xr[lane];
                v[i][1] = xr[lane + 32];
            } else {
                v[i][0] = v[i][1] = make_float4(0, 0, 0, 0);
            }
        }
        const float gw = -1.4426950408889634f * p.pw1_w1, gb = -1.4426950408889634f * p.pw1_b1;
        const float4 g0 = reinterpret_cast<const float4*>(p.ln_g)[lane], g1 = reinterpret_cast<const float4*>(p.ln_g)[lane + 32];
        const float4 b0 = reinterpret_cast<const float4*>(p.ln_b)[lane], b1 = reinterpret_cast<const float4*>(p.ln_b)[lane + 32];
#pragma unroll
        for (int i = 0; i < RB; ++i) {
            const int r = rb + i * NW, t = t0 - CM_HALO + r;
            if (r >= rows) break;                      // warp-uniform
            float4* urow = reinterpret_cast<float4*>(U + r * CM_D);
            if (t < 0 || t >= T) {
                urow[lane] = make_float4(0, 0, 0, 0);
                urow[lane + 32] = make_float4(0, 0, 0, 0);
                continue;
            }
            const float4 v0 = v[i][0], v1 = v[i][1];
            const float s = v0.x + v0.y + v0.z + v0.w + v1.x + v1.y + v1.z + v1.w;
            const float mean = warp_sum(s) / CM_D;
            float q = 0.f;
            {
                float a = v0.x - mean, b = v0.y - mean, c = v0.z - mean, d = v0.w - mean;
                q += a * a + b * b + c * c + d * d;
                a = v1.x - mean; b = v1.y - mean; c = v1.z - mean; d = v1.w - mean;
                q += a * a + b * b + c * c + d * d;
            }
            const float rstd = rsqrtf(warp_sum(q) / CM_D + 1e-5f);
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const float4 vv = hh ? v1 : v0, g = hh ? g1 : g0, bb = hh ? b1 : b0;
                float y[4] = {(vv.x - mean) * rstd * g.x + bb.x, (vv.y - mean) * rstd * g.y + bb.y,
                              (vv.z - mean) * rstd * g.z + bb.z, (vv.w - mean) * rstd * g.w + bb.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    // a * sigmoid(gt) = a / (1 + 2^(-gt log2 e)): the gate's affine map carries the -log2 e, one MUFU.EX2 and one
                    // MUFU.RCP (1 + 2^x is in [1, inf]: no range scaling needed, which div.approx would add)
                    const float a = y[e] * p.pw1_w0 + p.pw1_b0;
                    const float ge = fmaf(y[e], gw, gb);
                    float ex, rc;
                    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"(ge));
                    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(1.f + ex));
                    y[e] = a * rc;
                }
                urow[lane + 32 * hh] = make_float4(y[0], y[1], y[2], y[3]);
            }
        }
    }
    const int c = 2 * (threadIdx.x & (CM_D / 2 - 1)), part = threadIdx.x >> 7;     // channels c, c+1
    constexpr int NPART = THREADS / 128;
    float2 w[CM_TAPS];
    if (WSMEM) asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (WSMEM) {
#pragma unroll
        for (int k = 0; k < CM_TAPS; ++k) w[k] = make_float2(Wt[k * CM_WPITCH + c], Wt[k * CM_WPITCH + c + 1]);
    } else {
#pragma unroll
        for (int k = 0; k < CM_TAPS; ++k) w[k] = __ldg(reinterpret_cast<const float2*>(p.dw_wt + k * CM_D + c));
    }
    const float2 bias = *reinterpret_cast<const float2*>(p.dw_b + c);
    const float2 sc = *reinterpret_cast<const float2*>(p.bn_scale + c), sh = *reinterpret_cast<const float2*>(p.bn_shift + c);
    float* os = out + n * (long)T * CM_D;
    const int per = ((tile + NPART * CM_OUT - 1) / (NPART * CM_OUT)) * CM_OUT;       // multiple of CM_OUT
    const int t_end = min(tile, (part + 1) * per);
    for (int tt = part * per; tt < t_end; tt += CM_OUT) {
        float2 acc[CM_OUT], xv[CM_OUT];
#pragma unroll
        for (int j = 0; j < CM_OUT; ++j) {
            acc[j] = bias;
            // the residual rows are requested before the filter pass and consumed after it
            xv[j] = (tt + j < t_end) ? *reinterpret_cast<const float2*>(xs + (long)(t0 + tt + j) * CM_D + c) : make_float2(0.f, 0.f);
        }
#pragma unroll
        for (int kk = 0; kk < CM_TAPS + CM_OUT - 1; ++kk) {
            const int r = tt + kk;
            // rows past the tile's halo (last, partial pass) exist in the allocation and feed only outputs that are not stored
            const float2 u = *reinterpret_cast<const float2*>(U + r * CM_D + c);
#pragma unroll
            for (int j = 0; j < CM_OUT; ++j)
                if (kk - j >= 0 && kk - j < CM_TAPS) acc[j] = __ffma2_rn(w[kk - j], u, acc[j]);
        }
#pragma unroll
        for (int j = 0; j < CM_OUT; ++j) {
            if (tt + j < t_end) {
                const long idx = (long)(t0 + tt + j) * CM_D + c;
                float2 o;
                o.x = xv[j].x + (fmaxf(acc[j].x * sc.x + sh.x, 0.f) * p.pw2_w + p.pw2_b);
                o.y = xv[j].y + (fmaxf(acc[j].y * sc.y + sh.y, 0.f) * p.pw2_w + p.pw2_b);
                *reinterpret_cast<float2*>(os + idx) = o;
            }
        }
    }
}

// taps (d, 33) -> (33, d), once per model
__global__ void conv_taps_transpose_kernel(const float* __restrict__ w, float* __restrict__ wt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < CM_D * CM_TAPS) wt[(i % CM_TAPS) * CM_D + i / CM_TAPS] = w[i];
}
int conv_taps_transpose_launch(const float* w, float* wt, cudaStream_t st) {
    conv_taps_transpose_kernel<<<(CM_D * CM_TAPS + 255) / 256, 256, 0, st>>>(w, wt);
    return css_check_launch("conv_taps_transpose");
}

int conv_module_launch(const float* x, float* out, long N, int T, int d, const ConvModuleParams& p, cudaStream_t st) {
    CSS_REQUIRE(d == CM_D, "conv_module: d=%d unsupported (kernel is built for 256)", d);
    CSS_REQUIRE(p.ksize == CM_TAPS, "conv_module: taps must be centre-padded to %d (got %d)", CM_TAPS, p.ksize);
    CSS_REQUIRE(N > 0 && T > 0 && x != out, "conv_module: bad arguments");
    // two CTAs per SM on half-size tiles when the caller supplies the transposed taps (A/B: CSS_B200_CONV_SPLIT=0)
    static const bool split_env = !(getenv("CSS_B200_CONV_SPLIT") && getenv("CSS_B200_CONV_SPLIT")[0] == '0');
    const bool split = p.dw_wt != nullptr && split_env;
    const int max_rows = split ? CM_SPLIT_ROWS : CM_MAX_ROWS;
    const int max_tile = max_rows - 2 * CM_HALO;
    ConvPlan pl;
    pl.full_tiles = (T + max_tile - 1) / max_tile;
    pl.full_len = (T + pl.full_tiles - 1) / pl.full_tiles;
    pl.n_full = (int)N;
    pl.tail_tiles = pl.full_tiles;
    pl.tail_len = pl.full_len;
    const int slots = sm_count() * (split ? 2 : 1);            // CTAs per wave
    const int seq_slots = slots / pl.full_tiles;               // sequences per wave
    const int rem = seq_slots > 0 ? (int)(N % seq_slots) : 0;
    if (seq_slots > 0 && N > seq_slots && rem > 0 && T >= 16) {
        // cut the sequences of the last, partial wave into short tiles (at least 8 frames) that fill the idle SMs
        int tiles = slots / rem;
        if (tiles > T / 8) tiles = T / 8;
        if (tiles > pl.full_tiles) {
            pl.tail_len = (T + tiles - 1) / tiles;
            pl.tail_tiles = (T + pl.tail_len - 1) / pl.tail_len;
            pl.n_full = (int)N - rem;
        }
    }
    const long ctas = (long)pl.n_full * pl.full_tiles + (N - pl.n_full) * pl.tail_tiles;
    CSS_REQUIRE(ctas < (1L << 31), "conv_module: too many sequences (%ld)", N);
    {
        // the filter passes of a part read CM_OUT + 2 * CM_HALO rows from their first output row on: the last pass of the last
        // part ends at per * parts + 2 * CM_HALO
        const int parts = split ? 2 : 4;
        const int per = ((pl.full_len + parts * CM_OUT - 1) / (parts * CM_OUT)) * CM_OUT;
        pl.smem_rows = per * parts + 2 * CM_HALO;
        CSS_REQUIRE(pl.smem_rows <= max_rows, "conv_module: tile of %d frames does not fit shared memory", pl.full_len);
    }
    const size_t smem = sizeof(float) * ((size_t)pl.smem_rows * CM_D + (split ? 0 : CM_TAPS * CM_WPITCH));
    static CssPerDevice attr_set;
    if (attr_set.first()) {
        cudaFuncSetAttribute(conv_module_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (CM_MAX_ROWS * CM_D + CM_TAPS * CM_WPITCH) * 4);
        cudaFuncSetAttribute(conv_module_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, CM_SPLIT_ROWS * CM_D * 4);
    }
    css_prof_begin(2, 2.0 * (double)N * T * CM_D * 4, st);         // read x + write out
    if (split) conv_module_kernel<256><<<(unsigned)ctas, 256, smem, st>>>(x, out, T, pl, p);
    else conv_module_kernel<512><<<(unsigned)ctas, 512, smem, st>>>(x, out, T, pl, p);
    css_prof_end(st);
    return css_check_launch("conv_module");
}

// ---------------------------------------------------------------------------------------------
// Mask head layout: m[(n*T + t)][c*nb + f]  ->  masks[n][f][t][c]   (css/executor/separator.py:100-104)
// 32x32 (t, f) tiles through shared memory; reads coalesced along f, writes coalesced along (t, c).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
masks_layout_kernel(const float* __restrict__ m, int ldm, float* __restrict__ out, int T, int nb, int C) {
    __shared__ float tile[4][32][33];
    const long n = blockIdx.z;
    const int f0 = blockIdx.x * 32, t0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 8 rows per pass
    for (int c = 0; c < C; ++c)
        for (int r = ty; r < 32; r += 8) {
            const int t = t0 + r, f = f0 + tx;
            tile[c][r][tx] = (t < T && f < nb) ? m[(n * T + t) * (long)ldm + c * nb + f] : 0.f;
        }
    __syncthreads();
    // out row (n, f) is T*C contiguous floats; write the [t0, t0+32) x C slab of 32 f rows
    const int span = 32 * C;
    for (int r = ty; r < 32; r += 8) {
        const int f = f0 + r;
        if (f >= nb) continue;
        float* orow = out + ((n * nb + f) * (long)T + t0) * C;
        for (int e = tx; e < span; e += 32) {
            const int tt = e / C, c = e - tt * C;
            if (t0 + tt < T) orow[e] = tile[c][tt][r];
        }
    }
}

int masks_layout_launch(const float* m, int ldm, float* out, long N, int T, int nb, int C, cudaStream_t st) {
    CSS_REQUIRE(C >= 1 && C <= 4, "masks_layout: %d outputs unsupported (max 4)", C);
    CSS_REQUIRE(N > 0 && N <= 65535, "masks_layout: N=%ld out of range", N);
    dim3 grid((nb + 31) / 32, (T + 31) / 32, (unsigned)N);
    masks_layout_kernel<<<grid, 256, 0, st>>>(m, ldm, out, T, nb, C);
    return css_check_launch("masks_layout");
}

extern "C" int css_masks_layout(const float* m, int ldm, float* out, long N, int T, int nb, int C, cudaStream_t st) {
    return masks_layout_launch(m, ldm, out, N, T, nb, C, st);
}
