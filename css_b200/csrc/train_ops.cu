// Kernels of the PIT training step (css/trainer/iteration_types.py:34-63): the backward passes of the
// Conformer building blocks (css/models/conformer.py:188-362), train-mode BatchNorm, the mask-head product and the
// operand re-layouts (bf16 split planes and their transposes) that feed the tcgen05 GEMM with dY, W^T, dY^T and X^T.
#include "common.cuh"
#include "train.h"
#include <cuda_bf16.h>

typedef __nv_bfloat16 bf16;

__device__ __forceinline__ float block_sum_256x(float v, float* red) {   // any block size <= 1024, all threads call
    v = warp_sum(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float s = 0.f;
    if (warp == 0) {
        s = lane < nw ? red[lane] : 0.f;
        s = warp_sum(s);
    }
    return s;   // valid in warp 0
}

// ---------------------------------------------------------------------------------------------
// fp32 matrix -> bf16 split planes, transposed planes, column sums (bias gradients)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
split_t_kernel(const float* __restrict__ x, long rows, int cols, int ld, bf16* __restrict__ out, long out_ps, int ldp,
               bf16* __restrict__ outT, long outT_ps, long rows_p, int n_planes, float* __restrict__ colsum, float cs_scale) {
    __shared__ float tile[32][33];
    const int c0 = blockIdx.x * 32;
    const long r0 = (long)blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int i = ty; i < 32; i += 8) {
        const long r = r0 + i;
        const int c = c0 + tx;
        float v = (r < rows && c < cols) ? x[r * ld + c] : 0.f;
        tile[i][tx] = v;
        if (out && r < rows && c < ldp) {
            for (int p = 0; p < n_planes; ++p) {
                const bf16 h = __float2bfloat16_rn(v);
                v -= __bfloat162float(h);
                out[p * out_ps + r * ldp + c] = h;
            }
        }
    }
    __syncthreads();
    if (outT) {
        for (int i = ty; i < 32; i += 8) {
            const int c = c0 + i;
            const long r = r0 + tx;
            if (c < cols && r < rows_p) {
                float v = tile[tx][i];
                for (int p = 0; p < n_planes; ++p) {
                    const bf16 h = __float2bfloat16_rn(v);
                    v -= __bfloat162float(h);
                    outT[p * outT_ps + (long)c * rows_p + r] = h;
                }
            }
        }
    }
    if (colsum && ty == 0 && c0 + tx < cols) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 32; ++i) s += tile[i][tx];
        atomicAdd(colsum + c0 + tx, cs_scale * s);
    }
}

int split_t_launch(const float* x, long rows, int cols, int ld, const PlaneMat* out, const PlaneMat* outT, long rows_p,
                   float* colsum, float cs_scale, cudaStream_t st) {
    CSS_REQUIRE(rows > 0 && cols > 0 && ld >= cols, "split_t: bad shape");
    CSS_REQUIRE(!outT || (rows_p >= rows && outT->ld == rows_p), "split_t: transposed pitch %d != %ld", outT ? outT->ld : 0, rows_p);
    CSS_REQUIRE(!out || out->ld >= cols, "split_t: plane pitch too small");
    const int n_planes = out ? out->planes : (outT ? outT->planes : 0);
    CSS_REQUIRE(!(out && outT) || out->planes == outT->planes, "split_t: plane counts differ");
    const int width = out ? out->ld : cols;
    const long height = outT ? rows_p : rows;
    dim3 grid((width + 31) / 32, (unsigned)((height + 31) / 32));
    split_t_kernel<<<grid, 256, 0, st>>>(x, rows, cols, ld, out ? (bf16*)out->ptr : nullptr, out ? out->plane_stride : 0,
                                         out ? out->ld : 0, outT ? (bf16*)outT->ptr : nullptr, outT ? outT->plane_stride : 0,
                                         rows_p, n_planes, colsum, cs_scale);
    return css_check_launch("split_t");
}

__global__ void __launch_bounds__(256)
transpose_planes_kernel(const bf16* __restrict__ in, long in_ps, int ld, long rows, int cols, bf16* __restrict__ outT,
                        long out_ps, long rows_p) {
    __shared__ unsigned short tile[64][66];
    const int p = blockIdx.z;
    const int c0 = blockIdx.x * 64;
    const long r0 = (long)blockIdx.y * 64;
    const unsigned short* src = reinterpret_cast<const unsigned short*>(in) + p * in_ps;
    unsigned short* dst = reinterpret_cast<unsigned short*>(outT) + p * out_ps;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int i = ty; i < 64; i += 8) {
        const long r = r0 + i;
        const int c = c0 + 2 * tx;
        unsigned v = 0;
        if (r < rows && c < cols) v = *reinterpret_cast<const unsigned*>(src + r * ld + c);   // cols and ld are even
        tile[i][2 * tx] = (unsigned short)(v & 0xffff);
        tile[i][2 * tx + 1] = (unsigned short)(v >> 16);
    }
    __syncthreads();
    for (int i = ty; i < 64; i += 8) {
        const int c = c0 + i;
        const long r = r0 + 2 * tx;
        if (c < cols && r < rows_p) {
            const unsigned v = (unsigned)tile[2 * tx][i] | ((unsigned)tile[2 * tx + 1][i] << 16);
            *reinterpret_cast<unsigned*>(dst + (long)c * rows_p + r) = v;                        // rows_p is even
        }
    }
}

int transpose_planes_launch(const PlaneMat& in, long rows, int cols, const PlaneMat& outT, long rows_p, cudaStream_t st) {
    CSS_REQUIRE(rows > 0 && cols > 0 && (cols & 1) == 0 && (in.ld & 1) == 0 && in.ld >= cols, "transpose_planes: bad shape");
    CSS_REQUIRE(rows_p >= rows && (rows_p & 63) == 0 && outT.ld == rows_p && in.planes == outT.planes, "transpose_planes: bad output");
    dim3 grid((cols + 63) / 64, (unsigned)(rows_p / 64), in.planes);
    transpose_planes_kernel<<<grid, 256, 0, st>>>((const bf16*)in.ptr, in.plane_stride, in.ld, rows, cols, (bf16*)outT.ptr,
                                                  outT.plane_stride, rows_p);
    return css_check_launch("transpose_planes");
}

// ---------------------------------------------------------------------------------------------
// LayerNorm backward over d = 256 (one warp per row, 8 values per lane), optionally preceded by the backward of the
// conv module's scalar GLU.  dgamma/dbeta are reduced per CTA before the global atomics.
// ---------------------------------------------------------------------------------------------
template <bool GLU>
__global__ void __launch_bounds__(256)
ln_bwd_kernel(LnBwdArgs a, long M, float eps) {
    __shared__ float red[8][256];
    __shared__ float sred[32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float dg[8], db[8], gam[8], bet[8];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float4 g4 = reinterpret_cast<const float4*>(a.gamma)[lane + 32 * i];
        gam[4 * i] = g4.x; gam[4 * i + 1] = g4.y; gam[4 * i + 2] = g4.z; gam[4 * i + 3] = g4.w;
        if (GLU) {
            const float4 b4 = reinterpret_cast<const float4*>(a.beta)[lane + 32 * i];
            bet[4 * i] = b4.x; bet[4 * i + 1] = b4.y; bet[4 * i + 2] = b4.z; bet[4 * i + 3] = b4.w;
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) dg[e] = db[e] = 0.f;
    float w0 = 0.f, w1 = 0.f, b0 = 0.f, b1 = 0.f, aw0 = 0.f, aw1 = 0.f, ab0 = 0.f, ab1 = 0.f;
    if (GLU) { w0 = a.pw1_w[0]; w1 = a.pw1_w[1]; b0 = a.pw1_b[0]; b1 = a.pw1_b[1]; }
    for (long row = (long)blockIdx.x * 8 + warp; row < M; row += (long)gridDim.x * 8) {
        float v[8], dn[8];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const float4 x4 = reinterpret_cast<const float4*>(a.x + row * 256)[lane + 32 * i];
            const float4 d4 = reinterpret_cast<const float4*>(a.dn + row * 256)[lane + 32 * i];
            v[4 * i] = x4.x; v[4 * i + 1] = x4.y; v[4 * i + 2] = x4.z; v[4 * i + 3] = x4.w;
            dn[4 * i] = d4.x; dn[4 * i + 1] = d4.y; dn[4 * i + 2] = d4.z; dn[4 * i + 3] = d4.w;
            if (a.relu_out) {
                const float4 o4 = reinterpret_cast<const float4*>(a.relu_out + row * 256)[lane + 32 * i];
                if (!(o4.x > 0.f)) dn[4 * i] = 0.f;
                if (!(o4.y > 0.f)) dn[4 * i + 1] = 0.f;
                if (!(o4.z > 0.f)) dn[4 * i + 2] = 0.f;
                if (!(o4.w > 0.f)) dn[4 * i + 3] = 0.f;
            }
        }
        float s = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) s += v[e];
        const float mean = warp_sum(s) * (1.f / 256.f);
        float q = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) { v[e] -= mean; q += v[e] * v[e]; }
        const float rstd = rsqrtf(warp_sum(q) * (1.f / 256.f) + eps);
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float xh = v[e] * rstd;
            if (GLU) {
                const float n = xh * gam[e] + bet[e];
                const float av = n * w0 + b0, gt = n * w1 + b1;
                const float sg = 1.f / (1.f + expf(-gt));
                const float da = dn[e] * sg, dgt = dn[e] * av * sg * (1.f - sg);
                aw0 += da * n; ab0 += da; aw1 += dgt * n; ab1 += dgt;
                dn[e] = da * w0 + dgt * w1;
            }
            dg[e] += dn[e] * xh;
            db[e] += dn[e];
            const float dy = dn[e] * gam[e];
            s1 += dy;
            s2 += dy * xh;
            v[e] = xh;
            dn[e] = dy;
        }
        s1 = warp_sum(s1) * (1.f / 256.f);
        s2 = warp_sum(s2) * (1.f / 256.f);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            float4 o;
            o.x = rstd * (dn[4 * i] - s1 - v[4 * i] * s2);
            o.y = rstd * (dn[4 * i + 1] - s1 - v[4 * i + 1] * s2);
            o.z = rstd * (dn[4 * i + 2] - s1 - v[4 * i + 2] * s2);
            o.w = rstd * (dn[4 * i + 3] - s1 - v[4 * i + 3] * s2);
            float4* dp = reinterpret_cast<float4*>(a.dx + row * 256) + lane + 32 * i;
            if (a.accumulate) { const float4 p = *dp; o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w; }
            *dp = o;
        }
    }
    // CTA reduction of the parameter gradients: column c = 4*(lane + 32 i) + e
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int e = 0; e < 4; ++e) red[warp][4 * (lane + 32 * i) + e] = pass ? db[4 * i + e] : dg[4 * i + e];
        __syncthreads();
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
        atomicAdd((pass ? a.dbeta : a.dgamma) + threadIdx.x, t);
    }
    if (GLU) {
        float t;
        t = block_sum_256x(aw0, sred); if (threadIdx.x == 0) atomicAdd(a.dpw1_w, t);
        t = block_sum_256x(aw1, sred); if (threadIdx.x == 0) atomicAdd(a.dpw1_w + 1, t);
        t = block_sum_256x(ab0, sred); if (threadIdx.x == 0) atomicAdd(a.dpw1_b, t);
        t = block_sum_256x(ab1, sred); if (threadIdx.x == 0) atomicAdd(a.dpw1_b + 1, t);
    }
}

int ln_bwd_launch(const LnBwdArgs& a, long M, int d, float eps, bool glu, cudaStream_t st) {
    CSS_REQUIRE(d == 256, "ln_bwd: d=%d unsupported (kernel is built for 256)", d);
    CSS_REQUIRE(M > 0 && a.x && a.gamma && a.dn && a.dx && a.dgamma && a.dbeta, "ln_bwd: null argument");
    long grid = (M + 7) / 8;
    if (grid > 148 * 4) grid = 148 * 4;
    if (glu) ln_bwd_kernel<true><<<(unsigned)grid, 256, 0, st>>>(a, M, eps);
    else ln_bwd_kernel<false><<<(unsigned)grid, 256, 0, st>>>(a, M, eps);
    return css_check_launch("ln_bwd");
}

// ---------------------------------------------------------------------------------------------
// Conv module in training mode (css/models/conformer.py:289-303 with BatchNorm1d batch statistics).
// ---------------------------------------------------------------------------------------------
#define CT_D 256
#define CT_HALO 16
#define CT_TAPS 33

// LayerNorm + scalar GLU of one 256-wide row held 8 values per lane; returns u
__device__ __forceinline__ void ln_glu_row(const float* __restrict__ xr, const float* __restrict__ g, const float* __restrict__ b,
                                           float w0, float b0, float w1, float b1, int lane, float4* u) {
    const float4* x4 = reinterpret_cast<const float4*>(xr);
    float4 v[2] = {x4[lane], x4[lane + 32]};
    const float s = v[0].x + v[0].y + v[0].z + v[0].w + v[1].x + v[1].y + v[1].z + v[1].w;
    const float mean = warp_sum(s) / CT_D;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float a = v[i].x - mean, bb = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
        q += a * a + bb * bb + c * c + d * d;
    }
    const float rstd = rsqrtf(warp_sum(q) / CT_D + 1e-5f);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float4 gg = reinterpret_cast<const float4*>(g)[lane + 32 * i];
        const float4 bb = reinterpret_cast<const float4*>(b)[lane + 32 * i];
        float y[4] = {(v[i].x - mean) * rstd * gg.x + bb.x, (v[i].y - mean) * rstd * gg.y + bb.y,
                      (v[i].z - mean) * rstd * gg.z + bb.z, (v[i].w - mean) * rstd * gg.w + bb.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float a = y[e] * w0 + b0;
            const float gt = y[e] * w1 + b1;
            y[e] = a / (1.f + expf(-gt));
        }
        u[i] = make_float4(y[0], y[1], y[2], y[3]);
    }
}

// pre-BatchNorm activations c = dwconv(GLU(LN(x))) + bias, plus per-channel sum / sum of squares (fp64 atomics)
__global__ void __launch_bounds__(1024, 1)
conv_train_pre_kernel(const float* __restrict__ x, float* __restrict__ cpre, double* __restrict__ stats, int T, int tile_len,
                      ConvTrainParams p) {
    extern __shared__ float U[];   // [(tile + 2*CT_HALO)][CT_D]
    const long n = blockIdx.x;
    const int t0 = blockIdx.y * tile_len;
    const int tile = min(tile_len, T - t0);
    const int rows = tile + 2 * CT_HALO;
    const float* xs = x + n * (long)T * CT_D;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float w0 = p.pw1_w[0], w1 = p.pw1_w[1], b0 = p.pw1_b[0], b1 = p.pw1_b[1];
    for (int r = warp; r < rows; r += 32) {
        const int t = t0 - CT_HALO + r;
        float4* urow = reinterpret_cast<float4*>(U + r * CT_D);
        float4 u[2] = {make_float4(0, 0, 0, 0), make_float4(0, 0, 0, 0)};
        if (t >= 0 && t < T) ln_glu_row(xs + (long)t * CT_D, p.ln_g, p.ln_b, w0, b0, w1, b1, lane, u);
        urow[lane] = u[0];
        urow[lane + 32] = u[1];
    }
    __syncthreads();
    const int c = threadIdx.x & (CT_D - 1), part = threadIdx.x >> 8;
    float w[CT_TAPS];
#pragma unroll
    for (int k = 0; k < CT_TAPS; ++k) w[k] = p.dw_w[c * CT_TAPS + k];
    const float bias = p.dw_b[c];
    const int per = (tile + 3) / 4;
    const int t_end = min(tile, (part + 1) * per);
    float s1 = 0.f, s2 = 0.f;
    for (int tt = part * per; tt < t_end; ++tt) {
        float acc = bias;
#pragma unroll
        for (int k = 0; k < CT_TAPS; ++k) acc = fmaf(w[k], U[(tt + k) * CT_D + c], acc);
        cpre[(n * T + t0 + tt) * CT_D + c] = acc;
        s1 += acc;
        s2 += acc * acc;
    }
    atomicAdd(stats + c, (double)s1);
    atomicAdd(stats + CT_D + c, (double)s2);
}

__global__ void bn_finalize_kernel(const double* __restrict__ stats, double count, float* __restrict__ bnsave,
                                   float* __restrict__ running_mean, float* __restrict__ running_var) {
    const int c = threadIdx.x;
    const double mean = stats[c] / count;
    double var = stats[CT_D + c] / count - mean * mean;
    if (var < 0.0) var = 0.0;
    bnsave[c] = (float)mean;
    bnsave[CT_D + c] = (float)(1.0 / sqrt(var + 1e-5));
    if (running_mean) {
        running_mean[c] = 0.9f * running_mean[c] + 0.1f * (float)mean;
        running_var[c] = 0.9f * running_var[c] + 0.1f * (float)(var * count / (count - 1.0));
    }
}

__global__ void __launch_bounds__(256)
conv_train_apply_kernel(const float* __restrict__ x, const float* __restrict__ cpre, const float* __restrict__ bnsave,
                        float* __restrict__ out, long M, ConvTrainParams p, CssDrop dr, int use_drop) {
    const float w2 = p.pw2_w[0], b2 = p.pw2_b[0];
    const long total = M * (CT_D / 4);
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c4 = (int)(i & (CT_D / 4 - 1));
        const float4 cv = reinterpret_cast<const float4*>(cpre)[i];
        const float4 xv = reinterpret_cast<const float4*>(x)[i];
        const float4 mu = reinterpret_cast<const float4*>(bnsave)[c4];
        const float4 rs = reinterpret_cast<const float4*>(bnsave + CT_D)[c4];
        const float4 g = reinterpret_cast<const float4*>(p.bn_g)[c4];
        const float4 b = reinterpret_cast<const float4*>(p.bn_b)[c4];
        float4 o;
        if (use_drop) {
            const uint32_t e = (uint32_t)(i * 4);
            float br[4] = {fmaxf((cv.x - mu.x) * rs.x * g.x + b.x, 0.f) * w2 + b2, fmaxf((cv.y - mu.y) * rs.y * g.y + b.y, 0.f) * w2 + b2,
                           fmaxf((cv.z - mu.z) * rs.z * g.z + b.z, 0.f) * w2 + b2, fmaxf((cv.w - mu.w) * rs.w * g.w + b.w, 0.f) * w2 + b2};
#pragma unroll
            for (int k = 0; k < 4; ++k) br[k] = css_dropout_keep(dr, e + k) ? br[k] * dr.scale : 0.f;
            o = make_float4(xv.x + br[0], xv.y + br[1], xv.z + br[2], xv.w + br[3]);
        } else {
            o.x = xv.x + fmaxf((cv.x - mu.x) * rs.x * g.x + b.x, 0.f) * w2 + b2;
            o.y = xv.y + fmaxf((cv.y - mu.y) * rs.y * g.y + b.y, 0.f) * w2 + b2;
            o.z = xv.z + fmaxf((cv.z - mu.z) * rs.z * g.z + b.z, 0.f) * w2 + b2;
            o.w = xv.w + fmaxf((cv.w - mu.w) * rs.w * g.w + b.w, 0.f) * w2 + b2;
        }
        reinterpret_cast<float4*>(out)[i] = o;
    }
}

static int conv_tiles(int T, int max_tile, int* tile_len) {
    const int n_tiles = (T + max_tile - 1) / max_tile;
    *tile_len = (T + n_tiles - 1) / n_tiles;
    return n_tiles;
}

int conv_train_fwd_launch(const float* x, float* out, float* cpre, float* bnsave, double* stats, long N, int T,
                          const ConvTrainParams& p, const CssDrop* drop, cudaStream_t st) {
    CSS_REQUIRE(N > 0 && T > 0 && N * (long)T > 1 && x != out, "conv_train_fwd: bad arguments");
    int tile_len;
    const int n_tiles = conv_tiles(T, 188, &tile_len);
    const size_t smem = sizeof(float) * (size_t)(tile_len + 2 * CT_HALO) * CT_D;
    static CssPerDevice attr_set;
    if (attr_set.first()) {
        cudaFuncSetAttribute(conv_train_pre_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * CT_D * 4);
    }
    cudaMemsetAsync(stats, 0, 2 * CT_D * sizeof(double), st);
    dim3 grid((unsigned)N, n_tiles);
    conv_train_pre_kernel<<<grid, 1024, smem, st>>>(x, cpre, stats, T, tile_len, p);
    int rc = css_check_launch("conv_train_pre");
    if (rc) return rc;
    bn_finalize_kernel<<<1, CT_D, 0, st>>>(stats, (double)N * T, bnsave, p.running_mean, p.running_var);
    if ((rc = css_check_launch("bn_finalize"))) return rc;
    const long M = N * T;
    long g = (M * (CT_D / 4) + 255) / 256;
    if (g > 148 * 8) g = 148 * 8;
    CSS_REQUIRE(!drop || M * CT_D < (1L << 32), "conv_train_fwd: too many elements for the dropout index");
    conv_train_apply_kernel<<<(unsigned)g, 256, 0, st>>>(x, cpre, bnsave, out, M, p, drop ? *drop : CssDrop{0, 0, 1.f}, drop ? 1 : 0);
    return css_check_launch("conv_train_apply");
}

// per-channel sum dz and sum dz*zhat (dz = g * w2 * [z > 0]) + the gradients they are (BN beta / gamma) + pw_conv_2 grads
__global__ void __launch_bounds__(CT_D)
conv_bwd_stats_kernel(const float* __restrict__ g, const float* __restrict__ cpre, const float* __restrict__ bnsave, long M,
                      int rows_per_cta, float* __restrict__ bstats, ConvTrainParams p, ConvTrainGrads gr) {
    __shared__ float sred[32];
    const int c = threadIdx.x;
    const float mu = bnsave[c], rs = bnsave[CT_D + c], gam = p.bn_g[c], bet = p.bn_b[c], w2 = p.pw2_w[0];
    const long r0 = (long)blockIdx.x * rows_per_cta;
    const long r1 = min(M, r0 + rows_per_cta);
    float s1 = 0.f, s2 = 0.f, sw = 0.f, sb = 0.f;
    for (long r = r0; r < r1; ++r) {
        const float gv = g[r * CT_D + c];
        const float zh = (cpre[r * CT_D + c] - mu) * rs;
        const float z = zh * gam + bet;
        const float dz = z > 0.f ? gv * w2 : 0.f;
        s1 += dz;
        s2 += dz * zh;
        sw += gv * fmaxf(z, 0.f);
        sb += gv;
    }
    atomicAdd(bstats + c, s1);
    atomicAdd(bstats + CT_D + c, s2);
    atomicAdd(gr.bn_b + c, s1);
    atomicAdd(gr.bn_g + c, s2);
    float t;
    t = block_sum_256x(sw, sred); if (threadIdx.x == 0) atomicAdd(gr.pw2_w, t);
    t = block_sum_256x(sb, sred); if (threadIdx.x == 0) atomicAdd(gr.pw2_b, t);
}

#define CB_TILE 64
// BatchNorm backward (with the batch means from conv_bwd_stats) -> depthwise conv backward: du (to global), d taps, d bias
__global__ void __launch_bounds__(512, 1)
conv_bwd_dw_kernel(const float* __restrict__ x, const float* __restrict__ cpre, const float* __restrict__ bnsave,
                   const float* __restrict__ g, const float* __restrict__ bstats, float inv_count, float* __restrict__ du,
                   int T, ConvTrainParams p, ConvTrainGrads gr) {
    extern __shared__ float smem[];
    float* U = smem;                                         // [(CB_TILE + 32)][256]
    float* DC = smem + (CB_TILE + 2 * CT_HALO) * CT_D;       // [(CB_TILE + 32)][256]
    const long n = blockIdx.x;
    const int t0 = blockIdx.y * CB_TILE;
    const int tile = min(CB_TILE, T - t0);
    const int rows = tile + 2 * CT_HALO;
    const float* xs = x + n * (long)T * CT_D;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float w0 = p.pw1_w[0], w1 = p.pw1_w[1], b0 = p.pw1_b[0], b1 = p.pw1_b[1], w2 = p.pw2_w[0];
    for (int r = warp; r < rows; r += 16) {
        const int t = t0 - CT_HALO + r;
        float4 u[2] = {make_float4(0, 0, 0, 0), make_float4(0, 0, 0, 0)};
        float4 dc[2] = {make_float4(0, 0, 0, 0), make_float4(0, 0, 0, 0)};
        if (t >= 0 && t < T) {
            ln_glu_row(xs + (long)t * CT_D, p.ln_g, p.ln_b, w0, b0, w1, b1, lane, u);
            const long row = n * T + t;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int c4 = lane + 32 * i;
                const float4 gv = reinterpret_cast<const float4*>(g + row * CT_D)[c4];
                const float4 cv = reinterpret_cast<const float4*>(cpre + row * CT_D)[c4];
                const float4 mu = reinterpret_cast<const float4*>(bnsave)[c4];
                const float4 rs = reinterpret_cast<const float4*>(bnsave + CT_D)[c4];
                const float4 ga = reinterpret_cast<const float4*>(p.bn_g)[c4];
                const float4 be = reinterpret_cast<const float4*>(p.bn_b)[c4];
                const float4 m1 = reinterpret_cast<const float4*>(bstats)[c4];
                const float4 m2 = reinterpret_cast<const float4*>(bstats + CT_D)[c4];
                const float gg[4] = {gv.x, gv.y, gv.z, gv.w}, cc[4] = {cv.x, cv.y, cv.z, cv.w};
                const float mm[4] = {mu.x, mu.y, mu.z, mu.w}, rr[4] = {rs.x, rs.y, rs.z, rs.w};
                const float aa[4] = {ga.x, ga.y, ga.z, ga.w}, bb[4] = {be.x, be.y, be.z, be.w};
                const float a1[4] = {m1.x, m1.y, m1.z, m1.w}, a2[4] = {m2.x, m2.y, m2.z, m2.w};
                float o[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float zh = (cc[e] - mm[e]) * rr[e];
                    const float dz = (zh * aa[e] + bb[e]) > 0.f ? gg[e] * w2 : 0.f;
                    o[e] = aa[e] * rr[e] * (dz - a1[e] * inv_count - zh * a2[e] * inv_count);
                }
                dc[i] = make_float4(o[0], o[1], o[2], o[3]);
            }
        }
        reinterpret_cast<float4*>(U + r * CT_D)[lane] = u[0];
        reinterpret_cast<float4*>(U + r * CT_D)[lane + 32] = u[1];
        reinterpret_cast<float4*>(DC + r * CT_D)[lane] = dc[0];
        reinterpret_cast<float4*>(DC + r * CT_D)[lane + 32] = dc[1];
    }
    __syncthreads();
    const int c = threadIdx.x & (CT_D - 1), part = threadIdx.x >> 8;
    float w[CT_TAPS], dw[CT_TAPS];
#pragma unroll
    for (int k = 0; k < CT_TAPS; ++k) { w[k] = p.dw_w[c * CT_TAPS + k]; dw[k] = 0.f; }
    float dbias = 0.f;
    const int per = (tile + 1) / 2;
    const int t_end = min(tile, (part + 1) * per);
    for (int tt = part * per; tt < t_end; ++tt) {
        const int r = tt + CT_HALO;
        const float dcv = DC[r * CT_D + c];
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < CT_TAPS; ++k) {
            acc = fmaf(w[k], DC[(r + CT_HALO - k) * CT_D + c], acc);      // du[t] = sum_k w[k] dc[t + 16 - k]
            dw[k] = fmaf(dcv, U[(r + k - CT_HALO) * CT_D + c], dw[k]);    // dw[k] = sum_t dc[t] u[t + k - 16]
        }
        dbias += dcv;
        du[(n * T + t0 + tt) * CT_D + c] = acc;
    }
    __syncthreads();
    float* red = smem;     // [2][34][256]
#pragma unroll
    for (int k = 0; k < CT_TAPS; ++k) red[(part * 34 + k) * CT_D + c] = dw[k];
    red[(part * 34 + 33) * CT_D + c] = dbias;
    __syncthreads();
    for (int idx = threadIdx.x; idx < 34 * CT_D; idx += 512) {
        const int k = idx >> 8, cc = idx & 255;
        const float s = red[idx] + red[34 * CT_D + idx];
        if (k < CT_TAPS) atomicAdd(gr.dw_w + cc * CT_TAPS + k, s);
        else atomicAdd(gr.dw_b + cc, s);
    }
}

int conv_train_bwd_launch(const float* x, const float* cpre, const float* bnsave, float* g, const float* gbranch, float* du,
                          float* bstats, long N, int T, const ConvTrainParams& p, const ConvTrainGrads& gr, cudaStream_t st) {
    CSS_REQUIRE(N > 0 && T > 0, "conv_train_bwd: bad arguments");
    const long M = N * T;
    cudaMemsetAsync(bstats, 0, 2 * CT_D * sizeof(float), st);
    const int rows_per_cta = 32;
    conv_bwd_stats_kernel<<<(unsigned)((M + rows_per_cta - 1) / rows_per_cta), CT_D, 0, st>>>(gbranch, cpre, bnsave, M, rows_per_cta,
                                                                                             bstats, p, gr);
    int rc = css_check_launch("conv_bwd_stats");
    if (rc) return rc;
    const size_t smem = sizeof(float) * 2 * (size_t)(CB_TILE + 2 * CT_HALO) * CT_D;
    static CssPerDevice attr_set;
    if (attr_set.first()) {
        cudaFuncSetAttribute(conv_bwd_dw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    dim3 grid((unsigned)N, (T + CB_TILE - 1) / CB_TILE);
    conv_bwd_dw_kernel<<<grid, 512, smem, st>>>(x, cpre, bnsave, gbranch, bstats, 1.f / (float)M, du, T, p, gr);
    if ((rc = css_check_launch("conv_bwd_dw"))) return rc;
    LnBwdArgs a = {x, p.ln_g, p.ln_b, du, nullptr, g, 1, gr.ln_g, gr.ln_b, p.pw1_w, p.pw1_b, gr.pw1_w, gr.pw1_b};
    return ln_bwd_launch(a, M, CT_D, 1e-5f, true, st);
}

// ---------------------------------------------------------------------------------------------
// mask head product and its backward
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
ypred_kernel(const float* __restrict__ head, int ldh, const float* __restrict__ feat, int ldf, float* __restrict__ ypred, int T,
             int nb, int K) {
    const long row = blockIdx.x;           // n*T + t
    const long n = row / T;
    const int t = (int)(row - n * T);
    for (int i = threadIdx.x; i < K * nb; i += blockDim.x) {
        const int k = i / nb, f = i - k * nb;
        ypred[((n * K + k) * T + t) * nb + f] = head[row * ldh + i] * feat[row * ldf + f];
    }
}
int ypred_launch(const float* head, int ldh, const float* feat, int ldf, float* ypred, long N, int T, int nb, int K,
                 cudaStream_t st) {
    ypred_kernel<<<(unsigned)(N * T), 256, 0, st>>>(head, ldh, feat, ldf, ypred, T, nb, K);
    return css_check_launch("ypred");
}

__global__ void __launch_bounds__(256)
head_bwd_kernel(const float* __restrict__ head, int ldh, const float* __restrict__ feat, int ldf, const float* __restrict__ gy,
                float* __restrict__ dhead, int T, int nb, int K) {
    const long row = blockIdx.x;
    const long n = row / T;
    const int t = (int)(row - n * T);
    for (int i = threadIdx.x; i < ldh; i += blockDim.x) {
        float v = 0.f;
        if (i < K * nb) {
            const int k = i / nb, f = i - k * nb;
            if (head[row * ldh + i] > 0.f) v = gy[((n * K + k) * T + t) * nb + f] * feat[row * ldf + f];
        }
        dhead[row * ldh + i] = v;
    }
}
int head_bwd_launch(const float* head, int ldh, const float* feat, int ldf, const float* gy, float* dhead, long N, int T,
                    int nb, int K, cudaStream_t st) {
    head_bwd_kernel<<<(unsigned)(N * T), 256, 0, st>>>(head, ldh, feat, ldf, gy, dhead, T, nb, K);
    return css_check_launch("head_bwd");
}

// ---------------------------------------------------------------------------------------------
// optimiser: global gradient norm (clip_grad_norm_, css/trainer/iteration_types.py:54) and Adam with L2 weight decay
// (torch.optim.Adam, scripts/python/train.py:209-211), both over the flat parameter / gradient buffers
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sqnorm_kernel(const float* __restrict__ g, long n, double* __restrict__ out) {
    __shared__ double red[8];
    double acc = 0.0;
    const long n4 = n >> 2;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
        const float4 v = reinterpret_cast<const float4*>(g)[i];
        acc += (double)(v.x * v.x + v.y * v.y) + (double)(v.z * v.z + v.w * v.w);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)
        for (long i = n4 << 2; i < n; ++i) acc += (double)g[i] * g[i];
    acc = warp_sum_d(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += red[w];
        atomicAdd(out, s);
    }
}

extern "C" int css_grad_sqnorm(const float* g, long n, double* out, cudaStream_t st) {
    CSS_REQUIRE(g && out && n > 0 && ((uintptr_t)g & 15) == 0, "css_grad_sqnorm: bad arguments");
    cudaMemsetAsync(out, 0, sizeof(double), st);
    long grid = (n / 4 + 255) / 256;
    if (grid > 148 * 8) grid = 148 * 8;
    if (grid < 1) grid = 1;
    sqnorm_kernel<<<(unsigned)grid, 256, 0, st>>>(g, n, out);
    return css_check_launch("css_grad_sqnorm");
}

__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, long n, float lr,
            float beta1, float beta2, float eps, float wd, float bc1, float bc2_sqrt, float gscale, const double* __restrict__ sqnorm,
            float max_norm, float* __restrict__ norm_out) {
    float coef = gscale;
    if (sqnorm) {
        const float total = (float)sqrt(*sqnorm) * gscale;        // norm of the (already averaged) gradient
        if (norm_out && blockIdx.x == 0 && threadIdx.x == 0) *norm_out = total;
        if (max_norm > 0.f) coef *= fminf(1.f, max_norm / (total + 1e-6f));
    }
    const float step = lr / bc1;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const float pv = p[i];
        const float gv = g[i] * coef + wd * pv;
        const float mv = beta1 * m[i] + (1.f - beta1) * gv;
        const float vv = beta2 * v[i] + (1.f - beta2) * gv * gv;
        m[i] = mv;
        v[i] = vv;
        p[i] = pv - step * mv / (sqrtf(vv) / bc2_sqrt + eps);
    }
}

// One Adam step over flat buffers.  The gradient is first scaled by gscale (1/world after a sum all-reduce) and, when
// sqnorm (the squared norm of the unscaled gradient, css_grad_sqnorm) is given and max_norm > 0, clipped to max_norm
// exactly as torch.nn.utils.clip_grad_norm_ does.  step = 1-based update count.  norm_out (optional) receives the norm.
extern "C" int css_adam_step(float* p, const float* g, float* m, float* v, long n, float lr, float beta1, float beta2, float eps,
                             float weight_decay, long step, float gscale, const double* sqnorm, float max_norm, float* norm_out,
                             cudaStream_t st) {
    CSS_REQUIRE(p && g && m && v && n > 0 && step >= 1, "css_adam_step: bad arguments");
    const float bc1 = 1.f - powf(beta1, (float)step);
    const float bc2 = sqrtf(1.f - powf(beta2, (float)step));
    long grid = (n + 255) / 256;
    if (grid > 148 * 8) grid = 148 * 8;
    adam_kernel<<<(unsigned)grid, 256, 0, st>>>(p, g, m, v, n, lr, beta1, beta2, eps, weight_decay, bc1, bc2, gscale, sqnorm,
                                               max_norm, norm_out);
    return css_check_launch("css_adam_step");
}
