// Relative-position multi-head self-attention core on the warp-level tensor-core path
// (css/models/conformer.py:229-250: A = Q K^T, B[i][j] = q_i . pe[i-j], softmax((A+B)/sqrt(dk)), P V).
//
// One persistent CTA per SM loops over (sequence, head) items.  The 2T-1 relative-position rows of the
// embedding table stay resident in shared memory (bf16 hi/lo planes) for the whole launch; K and V of the
// current item are converted fp32 -> bf16 hi/lo planes on the way into shared memory.  Warp w owns query rows
// [16w, 16w+16): Q fragments live in registers, scores (16 x T) stay in registers, every product is the
// three-term split  hi*hi + hi*lo + lo*hi  with fp32 accumulation (mma.sync.m16n8k16.bf16), which keeps the
// result at fp32-class accuracy.  The relative term needs R[i][i-j]: R is produced 8 columns at a time in the
// "reversed" column order u = j + (i0+15-i), parked in a small per-warp ring in shared memory and read back
// skewed by the row, so no 150x150x64 gather is ever materialised (the reference builds pos_k (T,T,64)).
//
// T <= 160 (the executor always runs 150-frame chunks); other lengths use the SIMT kernel in conformer_ops.cu.
#include "common.cuh"
#include "conformer.h"
#include <cuda_bf16.h>

namespace {

constexpr int DK = 64;
constexpr int TMAX = 160;
constexpr int NWARP = TMAX / 16;            // 10 warps, one 16-row query tile each
constexpr int ROWB = 144;                   // bytes per smem row: 64 bf16 + 16 B pad (conflict-free ldmatrix)
constexpr int PE_ROWS = 2 * TMAX;           // >= 2T-1
constexpr int PLANE_KV = TMAX * ROWB;       // one bf16 plane of K or V
constexpr int PLANE_PE = PE_ROWS * ROWB;
constexpr int RING_BYTES = 4 * 16 * 8 * 4;  // 4 slots of a 16 x 8 fp32 tile per warp
constexpr int SMEM_BYTES = 2 * PLANE_PE + 4 * PLANE_KV + NWARP * RING_BYTES;

__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm4t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
// (x, y) -> packed bf16 hi pair and lo (residue) pair; element x in the low half
__device__ __forceinline__ void split2(float x, float y, uint32_t& hi, uint32_t& lo) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(x, y);
    const __nv_bfloat162 l = __floats2bfloat162_rn(x - __bfloat162float(h.x), y - __bfloat162float(h.y));
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}
// fp32 row (64 values) -> hi/lo planes in smem; 16 lanes x float4 cover a row
__device__ __forceinline__ void store_row_split(uint8_t* hi_plane, uint8_t* lo_plane, int row, int c4, float4 v) {
    uint32_t h0, l0, h1, l1;
    split2(v.x, v.y, h0, l0);
    split2(v.z, v.w, h1, l1);
    *reinterpret_cast<uint2*>(hi_plane + row * ROWB + c4 * 8) = make_uint2(h0, h1);
    *reinterpret_cast<uint2*>(lo_plane + row * ROWB + c4 * 8) = make_uint2(l0, l1);
}

// three-term product of one 16x16 A tile (hi/lo) with one 16x8 B tile (hi/lo)
__device__ __forceinline__ void mma3(float* c, const uint32_t* ah, const uint32_t* al, uint32_t bh0, uint32_t bh1,
                                     uint32_t bl0, uint32_t bl1) {
    mma16816(c, al, bh0, bh1);
    mma16816(c, ah, bl0, bl1);
    mma16816(c, ah, bh0, bh1);
}

__global__ void __launch_bounds__(NWARP * 32, 1)
attention_tc_kernel(const float* __restrict__ qkv, const float* __restrict__ pe, int pe_rows, float* __restrict__ out,
                    __nv_bfloat16* __restrict__ planes, long plane_stride, int n_planes, long n_items, int T, int H) {
    extern __shared__ __align__(16) uint8_t sm[];
    uint8_t* pe_hi = sm;
    uint8_t* pe_lo = pe_hi + PLANE_PE;
    uint8_t* k_hi = pe_lo + PLANE_PE;
    uint8_t* k_lo = k_hi + PLANE_KV;
    uint8_t* v_hi = k_lo + PLANE_KV;
    uint8_t* v_lo = v_hi + PLANE_KV;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* ring = reinterpret_cast<float*>(v_lo + PLANE_KV) + warp * (RING_BYTES / 4);
    const int g = lane >> 2, q = lane & 3;
    const int ld = 3 * H * DK;
    const int maxlen = pe_rows / 2;
    const bool use_pe = pe != nullptr;
    const int TK = (T + 15) & ~15;                 // keys padded to the PV k-step
    const int n_stiles = (T + 7) >> 3;             // score n-tiles (8 keys each)

    // ---- relative-position rows, resident for the whole launch: smem row rho <-> r = rho - (T-1) ----
    if (use_pe) {
        const int nrel = 2 * T - 1;
        for (int i = threadIdx.x; i < PE_ROWS * 16; i += blockDim.x) {
            const int rho = i >> 4, c4 = i & 15;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (rho < nrel) {
                int r = rho - (T - 1);
                r = max(-maxlen, min(maxlen - 1, r)) + maxlen;
                v = *reinterpret_cast<const float4*>(pe + (long)r * DK + c4 * 4);
            }
            store_row_split(pe_hi, pe_lo, rho, c4, v);
        }
    }
    const uint32_t pe_hi_s = (uint32_t)__cvta_generic_to_shared(pe_hi), pe_lo_s = (uint32_t)__cvta_generic_to_shared(pe_lo);
    const uint32_t k_hi_s = (uint32_t)__cvta_generic_to_shared(k_hi), k_lo_s = (uint32_t)__cvta_generic_to_shared(k_lo);
    const uint32_t v_hi_s = (uint32_t)__cvta_generic_to_shared(v_hi), v_lo_s = (uint32_t)__cvta_generic_to_shared(v_lo);
    const int i0 = warp * 16;

    for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const long n = item / H;
        const int h = (int)(item - n * H);
        const float* base = qkv + n * (long)T * ld;
        __syncthreads();                           // previous item's K/V fully consumed (and PE rows written)
        for (int i = threadIdx.x; i < TK * 16; i += blockDim.x) {
            const int j = i >> 4, c4 = i & 15;
            float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
            if (j < T) {
                kv = *reinterpret_cast<const float4*>(base + (long)j * ld + (H + h) * DK + c4 * 4);
                vv = *reinterpret_cast<const float4*>(base + (long)j * ld + (2 * H + h) * DK + c4 * 4);
            }
            store_row_split(k_hi, k_lo, j, c4, kv);
            store_row_split(v_hi, v_lo, j, c4, vv);
        }
        __syncthreads();
        if (i0 >= T) continue;                     // warp-uniform; the barriers above are still reached next round

        // ---- Q fragments (hi/lo) of rows i0+g and i0+g+8 ----
        uint32_t qh[4][4], ql[4][4];
        {
            const int r0 = i0 + g, r1 = i0 + g + 8;
            const float* q0 = base + (long)r0 * ld + h * DK;
            const float* q1 = base + (long)r1 * ld + h * DK;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const int c = 16 * ks + 2 * q;
                float2 a0 = make_float2(0.f, 0.f), a1 = a0, a2 = a0, a3 = a0;
                if (r0 < T) { a0 = *reinterpret_cast<const float2*>(q0 + c); a2 = *reinterpret_cast<const float2*>(q0 + c + 8); }
                if (r1 < T) { a1 = *reinterpret_cast<const float2*>(q1 + c); a3 = *reinterpret_cast<const float2*>(q1 + c + 8); }
                split2(a0.x, a0.y, qh[ks][0], ql[ks][0]);
                split2(a1.x, a1.y, qh[ks][1], ql[ks][1]);
                split2(a2.x, a2.y, qh[ks][2], ql[ks][2]);
                split2(a3.x, a3.y, qh[ks][3], ql[ks][3]);
            }
        }
        // R tile v: columns u = 8v .. 8v+7, u = j + (15 - l)  <->  relative position r = i0 + 15 - u
        auto rel_tile = [&](int v) {
            float c[4] = {0.f, 0.f, 0.f, 0.f};
            int rho = (i0 + 15 - (8 * v + (lane & 7))) + (T - 1);
            rho = max(0, min(2 * T - 2, rho));
            const uint32_t off = rho * ROWB + (lane >> 3) * 16;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                uint32_t bh[4], bl[4];
                ldsm4(pe_hi_s + off + half * 64, bh[0], bh[1], bh[2], bh[3]);
                ldsm4(pe_lo_s + off + half * 64, bl[0], bl[1], bl[2], bl[3]);
                mma3(c, qh[2 * half], ql[2 * half], bh[0], bh[1], bl[0], bl[1]);
                mma3(c, qh[2 * half + 1], ql[2 * half + 1], bh[2], bh[3], bl[2], bl[3]);
            }
            float* slot = ring + (v & 3) * 128;
            *reinterpret_cast<float2*>(slot + g * 8 + 2 * q) = make_float2(c[0], c[1]);
            *reinterpret_cast<float2*>(slot + (g + 8) * 8 + 2 * q) = make_float2(c[2], c[3]);
        };

        float s[TMAX / 8][4];
        if (use_pe) { rel_tile(0); rel_tile(1); }
#pragma unroll
        for (int t = 0; t < TMAX / 8; ++t) {
            if (t < n_stiles) {
                if (use_pe) rel_tile(t + 2);
                __syncwarp();
                float* c = s[t];
                c[0] = c[1] = c[2] = c[3] = 0.f;
                const uint32_t off = (8 * t + (lane & 7)) * ROWB + (lane >> 3) * 16;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    uint32_t bh[4], bl[4];
                    ldsm4(k_hi_s + off + half * 64, bh[0], bh[1], bh[2], bh[3]);
                    ldsm4(k_lo_s + off + half * 64, bl[0], bl[1], bl[2], bl[3]);
                    mma3(c, qh[2 * half], ql[2 * half], bh[0], bh[1], bl[0], bl[1]);
                    mma3(c, qh[2 * half + 1], ql[2 * half + 1], bh[2], bh[3], bl[2], bl[3]);
                }
                if (use_pe) {
                    // element (l, j): u = j + 15 - l ;  l = g (+8), j = 8t + 2q (+1)
                    const int u0 = 8 * t + 2 * q + 15 - g;           // row g
                    const int u1 = u0 - 8;                           // row g + 8
                    c[0] += ring[((u0 >> 3) & 3) * 128 + g * 8 + (u0 & 7)];
                    c[1] += ring[(((u0 + 1) >> 3) & 3) * 128 + g * 8 + ((u0 + 1) & 7)];
                    c[2] += ring[((u1 >> 3) & 3) * 128 + (g + 8) * 8 + (u1 & 7)];
                    c[3] += ring[(((u1 + 1) >> 3) & 3) * 128 + (g + 8) * 8 + ((u1 + 1) & 7)];
                }
            }
        }
        // ---- softmax over keys (rows g and g+8 live in the 4 lanes of a quad) ----
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int t = 0; t < TMAX / 8; ++t) {
            if (t < n_stiles) {
                const int j = 8 * t + 2 * q;
                s[t][0] = (j < T) ? s[t][0] * 0.125f : -INFINITY;
                s[t][1] = (j + 1 < T) ? s[t][1] * 0.125f : -INFINITY;
                s[t][2] = (j < T) ? s[t][2] * 0.125f : -INFINITY;
                s[t][3] = (j + 1 < T) ? s[t][3] * 0.125f : -INFINITY;
                mx0 = fmaxf(mx0, fmaxf(s[t][0], s[t][1]));
                mx1 = fmaxf(mx1, fmaxf(s[t][2], s[t][3]));
            }
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
        for (int t = 0; t < TMAX / 8; ++t) {
            if (t < n_stiles) {
                s[t][0] = expf(s[t][0] - mx0); s[t][1] = expf(s[t][1] - mx0);
                s[t][2] = expf(s[t][2] - mx1); s[t][3] = expf(s[t][3] - mx1);
                sum0 += s[t][0] + s[t][1];
                sum1 += s[t][2] + s[t][3];
            } else {
                s[t][0] = s[t][1] = s[t][2] = s[t][3] = 0.f;
            }
        }
        sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1); sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
        sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1); sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
        const float inv0 = 1.f / sum0, inv1 = 1.f / sum1;

        // ---- O = P V : probabilities become A fragments straight from the score registers ----
        float o[8][4];
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
#pragma unroll
        for (int kt = 0; kt < TMAX / 16; ++kt) {
            if (16 * kt < T) {
                uint32_t ph[4], pl[4];
                split2(s[2 * kt][0] * inv0, s[2 * kt][1] * inv0, ph[0], pl[0]);
                split2(s[2 * kt][2] * inv1, s[2 * kt][3] * inv1, ph[1], pl[1]);
                split2(s[2 * kt + 1][0] * inv0, s[2 * kt + 1][1] * inv0, ph[2], pl[2]);
                split2(s[2 * kt + 1][2] * inv1, s[2 * kt + 1][3] * inv1, ph[3], pl[3]);
                const uint32_t off = (16 * kt + (lane & 7) + 8 * ((lane >> 3) & 1)) * ROWB + (lane >> 4) * 16;
#pragma unroll
                for (int np = 0; np < 4; ++np) {           // two 8-wide output tiles per ldmatrix.x4
                    uint32_t bh[4], bl[4];
                    ldsm4t(v_hi_s + off + np * 32, bh[0], bh[1], bh[2], bh[3]);
                    ldsm4t(v_lo_s + off + np * 32, bl[0], bl[1], bl[2], bl[3]);
                    mma3(o[2 * np], ph, pl, bh[0], bh[1], bl[0], bl[1]);
                    mma3(o[2 * np + 1], ph, pl, bh[2], bh[3], bl[2], bl[3]);
                }
            }
        }
        // ---- store rows i0+g, i0+g+8: fp32 and/or bf16 split planes (the out-projection GEMM's A operand) ----
        const int dm = H * DK;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int r = i0 + g + 8 * half;
            if (r >= T) continue;
            const long row = n * T + r;
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                const int col = h * DK + 8 * nt + 2 * q;
                float x = o[nt][2 * half], y = o[nt][2 * half + 1];
                if (out) *reinterpret_cast<float2*>(out + row * dm + col) = make_float2(x, y);
                for (int p = 0; p < n_planes; ++p) {
                    const __nv_bfloat162 hb = __floats2bfloat162_rn(x, y);
                    x -= __bfloat162float(hb.x);
                    y -= __bfloat162float(hb.y);
                    *reinterpret_cast<__nv_bfloat162*>(planes + p * plane_stride + row * dm + col) = hb;
                }
            }
        }
    }
}

}  // namespace

bool attention_tc_supported(int T, int dk) { return dk == DK && T >= 1 && T <= TMAX; }

// out (fp32, optional) and/or bf16 planes (pitch H*64, optional)
int attention_tc_launch(const float* qkv, const float* pe, int pe_rows, float* out, void* planes, long plane_stride,
                        int n_planes, long N, int T, int H, int dk, cudaStream_t st) {
    CSS_REQUIRE(attention_tc_supported(T, dk), "attention_tc: T=%d / dk=%d unsupported", T, dk);
    CSS_REQUIRE(N > 0 && H > 0 && (out || (planes && n_planes > 0)), "attention_tc: bad arguments");
    static bool attr = false;
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        CSS_REQUIRE(e == cudaSuccess, "attention_tc: cannot reserve %d B of shared memory: %s", SMEM_BYTES, cudaGetErrorString(e));
        attr = true;
    }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long items = N * H;
    const int grid = (int)(items < sms ? items : sms);
    css_prof_begin(1, 6.0 * (double)N * H * T * T * DK, st);   // QK^T + Q.PE^T + PV (algorithmic)
    attention_tc_kernel<<<grid, NWARP * 32, SMEM_BYTES, st>>>(qkv, pe, pe_rows, out, static_cast<__nv_bfloat16*>(planes),
                                                              plane_stride, planes ? n_planes : 0, items, T, H);
    css_prof_end(st);
    return css_check_launch("attention_tc");
}

extern "C" int css_relpos_attention_tc(const float* qkv, const float* pe, int pe_rows, float* out, long N, int T, int H, int dk,
                                       cudaStream_t st) {
    return attention_tc_launch(qkv, pe, pe_rows, out, nullptr, 0, 0, N, T, H, dk, st);
}
