// Relative-position multi-head self-attention core on the warp-level tensor-core path
// (css/models/conformer.py:229-250: A = Q K^T, B[i][j] = q_i . pe[i-j], softmax((A+B)/sqrt(dk)), P V).
//
// One persistent CTA per SM loops over (sequence, head) items.  Q, K, V arrive as bf16 split planes (written by the
// QKV GEMM's epilogue), so the K/V tiles of an item are plain cp.async copies and the Q fragments plain 32-bit
// loads; the 2T-1 relative-position rows of the embedding table stay resident in shared memory (hi/lo planes)
// for the whole launch.  Warp w owns query rows
// [16w, 16w+16): Q fragments live in registers, scores (16 x T) stay in registers, every product is the
// three-term split  hi*hi + hi*lo + lo*hi  with fp32 accumulation (mma.sync.m16n8k16.bf16), which keeps the
// result at fp32-class accuracy.  The relative term needs R[i][i-j]: R is produced 8 columns at a time in the
// "reversed" column order u = j + (i0+15-i), parked in a small per-warp ring in shared memory and read back
// skewed by the row, so no 150x150x64 gather is ever materialised (the reference builds pos_k (T,T,64)).
//
// T <= 160 (the executor always runs 150-frame chunks); other lengths use the SIMT kernel in conformer_ops.cu.
#include "common.cuh"
#include "conformer.h"
#include "gemm.h"
#include <cuda_bf16.h>

namespace {

constexpr int DK = 64;
constexpr int TMAX = 160;
constexpr int NWARP = TMAX / 16;            // 10 warps, one 16-row query tile each
constexpr int ROWB = 144;                   // bytes per smem row: 64 bf16 + 16 B pad (conflict-free ldmatrix)
constexpr int PE_ROWS = 2 * TMAX;           // >= 2T-1
constexpr int PLANE_KV = TMAX * ROWB;       // one bf16 plane of K or V
constexpr int PLANE_PE = PE_ROWS * ROWB;
constexpr int SLOT_BYTES = 16 * 8 * 4;      // one 16 x 8 fp32 tile of the relative term
constexpr int RING_BYTES = 8 * SLOT_BYTES;  // 8 slots per warp, 4096-aligned so the read pointers wrap with one AND
constexpr int SMEM_BYTES = 2 * PLANE_PE + 4 * PLANE_KV + NWARP * RING_BYTES + 4096;

__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm(
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
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ float lds_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sts_f32x2(uint32_t addr, float x, float y) {
    asm volatile("st.shared.v2.f32 [%0], {%1,%2};" ::"r"(addr), "f"(x), "f"(y) : "memory");
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

struct AttnArgs {
    const __nv_bfloat16* qkv;   // [planes][M][3*H*64]
    long qkv_plane_stride;
    const __nv_bfloat16* pe;    // [planes][pe_rows][64] or null
    long pe_plane_stride;
    int pe_rows;
    float* out;                 // (M, H*64) fp32 or null
    __nv_bfloat16* planes;      // [n_planes][M][H*64] or null
    long plane_stride;
    int n_planes;
    long n_items;
    int T, H;
};

// one 16 x 8 tile, 64-deep contraction: lo[] += Al*Bh + Ah*Bl (SPLIT only), hi[] += Ah*Bh.  The two correction
// terms are ~2^-8 of the main one, so they share an accumulator; the caller adds lo into hi once per tile.
template <bool SPLIT>
__device__ __forceinline__ void tile_k64(float* lo, float* hi, const uint32_t (*ah)[4], const uint32_t (*al)[4], uint32_t hi_s,
                                         uint32_t lo_s, uint32_t off) {
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        uint32_t bh[4], bl[4];
        ldsm4(hi_s + off + half * 64, bh[0], bh[1], bh[2], bh[3]);
        if (SPLIT) ldsm4(lo_s + off + half * 64, bl[0], bl[1], bl[2], bl[3]);
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
            const int ks = 2 * half + kk;
            if (SPLIT) {
                mma16816(lo, al[ks], bh[2 * kk], bh[2 * kk + 1]);
                mma16816(lo, ah[ks], bl[2 * kk], bl[2 * kk + 1]);
            }
            mma16816(hi, ah[ks], bh[2 * kk], bh[2 * kk + 1]);
        }
    }
}

// NT: score tiles of 8 keys computed per query tile (T <= 8 NT).  The tile loops have compile-time trip counts and no
// run-time branches, so ldmatrix / mma of neighbouring tiles overlap; keys in [T, 8 NT) read zeroed rows and are masked.
template <bool SPLIT, bool PE, int NT>
__global__ void __launch_bounds__(NWARP * 32, 1)
attention_tc_kernel(const AttnArgs a) {
    extern __shared__ __align__(16) uint8_t sm[];
    const int T = a.T, H = a.H;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t sm_s = (uint32_t)__cvta_generic_to_shared(sm);
    const uint32_t pe_hi_s = sm_s, pe_lo_s = pe_hi_s + PLANE_PE;
    const uint32_t k_hi_s = pe_lo_s + PLANE_PE, k_lo_s = k_hi_s + PLANE_KV;
    const uint32_t v_hi_s = k_lo_s + PLANE_KV, v_lo_s = v_hi_s + PLANE_KV;
    const uint32_t ring_s = ((v_lo_s + PLANE_KV + 4095u) & ~4095u) + warp * RING_BYTES;
    const int g = lane >> 2, q = lane & 3;
    const int ld = 3 * H * DK;
    constexpr int NPL = SPLIT ? 2 : 1;
    constexpr int NKT = (NT + 1) / 2;              // PV k-steps of 16 keys

    // K (kv = 0) or V (kv = 1) planes of one (sequence, head) item -> shared memory; 40 rows per pass of the CTA
    auto load_kv = [&](long item, int kv) {
        const long n = item / H;
        const int h = (int)(item - n * H);
        const int j0 = threadIdx.x >> 3, c8 = threadIdx.x & 7;
#pragma unroll
        for (int p = 0; p < NPL; ++p) {
            const __nv_bfloat16* src = a.qkv + p * a.qkv_plane_stride + (n * T + j0) * (long)ld + ((kv ? 2 : 1) * H + h) * DK + c8 * 8;
            uint32_t dst = (kv ? (p ? v_lo_s : v_hi_s) : (p ? k_lo_s : k_hi_s)) + j0 * ROWB + c8 * 16;
            for (int j = j0; j < T; j += NWARP * 4) {
                cp_async16(dst, src);
                src += (long)(NWARP * 4) * ld;
                dst += NWARP * 4 * ROWB;
            }
        }
        cp_async_commit();
    };

    // ---- one-off: relative-position rows (smem row rho <-> r = rho - (T-1)), zero rows [T, TMAX) of K and V ----
    if (PE) {
        const int maxlen = a.pe_rows / 2;
        const int nrel = 2 * T - 1;
        for (int i = threadIdx.x; i < nrel * 8 * NPL; i += blockDim.x) {
            const int p = i / (nrel * 8), rem = i - p * (nrel * 8);
            const int rho = rem >> 3, c8 = rem & 7;
            int r = rho - (T - 1);
            r = max(-maxlen, min(maxlen - 1, r)) + maxlen;
            cp_async16((p ? pe_lo_s : pe_hi_s) + rho * ROWB + c8 * 16, a.pe + p * a.pe_plane_stride + (long)r * DK + c8 * 8);
        }
    }
    for (int i = threadIdx.x; i < (TMAX - T) * 8 * 4; i += blockDim.x) {
        const int pl = i / ((TMAX - T) * 8), rem = i - pl * ((TMAX - T) * 8);
        *reinterpret_cast<uint4*>(sm + 2 * PLANE_PE + pl * PLANE_KV + (T + (rem >> 3)) * ROWB + (rem & 7) * 16) = make_uint4(0, 0, 0, 0);
    }
    if (blockIdx.x < a.n_items) {
        load_kv(blockIdx.x, 0);                    // group: PE rows + K of the first item
        load_kv(blockIdx.x, 1);
    }
    const int i0 = warp * 16;
    const bool active = i0 < T;                    // warp-uniform; idle warps still reach every barrier
    const float kscale = 0.125f * 1.4426950408889634f;        // 1/sqrt(64) * log2(e)
    // lane constants of the relative term.  R tile v holds columns u = 8v .. 8v+7 with u = j + (15 - l), i.e. relative
    // position r = i0 + 15 - u; element (l, j) of score tile t sits at u = 8t + (j - 8t) + 15 - l.
    const int rho0 = i0 + 15 - (lane & 7) + (T - 1);
    const uint32_t rel_lane = (uint32_t)((lane >> 3) * 16);
    const uint32_t st_s = ring_s + (uint32_t)(g * 8 + 2 * q) * 4;
    uint32_t rd0[4];                               // ring read pointers of the lane's 4 elements at tile 0
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const int c = 2 * q + 15 - g - ((e >> 1) ? 8 : 0) + (e & 1);
        rd0[e] = ring_s + (uint32_t)((c >> 3) * SLOT_BYTES + ((g + ((e >> 1) ? 8 : 0)) * 8 + (c & 7)) * 4);
    }

    // Q fragments (hi/lo) of rows i0+g and i0+g+8 of one item: plain 32-bit loads from the planes.  The next item's are
    // requested right after PV, so they land under the output stores.
    uint32_t qh[4][4], ql[4][4];
    auto load_q = [&](long item) {
        const long n = item / H;
        const int h = (int)(item - n * H);
        const __nv_bfloat16* base = a.qkv + n * (long)T * ld;
        const int r0 = i0 + g, r1 = i0 + g + 8;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
            const int c = h * DK + 16 * ks + 2 * q;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int r = (e & 1) ? r1 : r0;
                const long off = (long)r * ld + c + ((e >> 1) ? 8 : 0);
                qh[ks][e] = (r < T) ? *reinterpret_cast<const uint32_t*>(base + off) : 0u;
                ql[ks][e] = (SPLIT && r < T) ? *reinterpret_cast<const uint32_t*>(base + a.qkv_plane_stride + off) : 0u;
            }
        }
    };
    if (blockIdx.x < a.n_items) load_q(blockIdx.x);

    for (long item = blockIdx.x; item < a.n_items; item += gridDim.x) {
        const long n = item / H;
        const int h = (int)(item - n * H);
        const long next = item + gridDim.x;

        cp_async_wait<1>();                        // K of this item (and the PE rows) landed; V may still be in flight
        __syncthreads();

        float s[2 * NKT][4];
        if (active) {
            auto rel_tile = [&](int v) {           // R tile v -> ring slot v & 7
                float lo[4] = {}, hi[4] = {};
                const int rho = max(0, min(2 * T - 2, rho0 - 8 * v));
                tile_k64<SPLIT>(lo, hi, qh, ql, pe_hi_s, pe_lo_s, (uint32_t)(rho * ROWB) + rel_lane);
                const uint32_t dst = st_s + (v & 7) * SLOT_BYTES;
                sts_f32x2(dst, hi[0] + lo[0], hi[1] + lo[1]);
                sts_f32x2(dst + 8 * 8 * 4, hi[2] + lo[2], hi[3] + lo[3]);
            };
            uint32_t rd[4] = {rd0[0], rd0[1], rd0[2], rd0[3]};
            if (PE) {
                rel_tile(0);
                rel_tile(1);
            }
            const uint32_t k_lane = (uint32_t)((lane & 7) * ROWB) + rel_lane;
#pragma unroll
            for (int t2 = 0; t2 < NKT; ++t2) {
                float lo[2][4] = {}, hi[2][4] = {};
#pragma unroll
                for (int b = 0; b < 2; ++b) {
                    const int t = 2 * t2 + b;
                    if (t < NT) tile_k64<SPLIT>(lo[b], hi[b], qh, ql, k_hi_s, k_lo_s, k_lane + (uint32_t)(8 * t * ROWB));
                    if (PE && t + 2 <= NT + 1) rel_tile(t + 2);
                }
                if (PE) __syncwarp();
#pragma unroll
                for (int b = 0; b < 2; ++b) {
                    const int t = 2 * t2 + b;
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        s[t][e] = hi[b][e] + lo[b][e];
                        if (PE && t < NT) s[t][e] += lds_f32(rd[e]);
                        if (PE) rd[e] = ((rd[e] + SLOT_BYTES) & (RING_BYTES - 1)) | ring_s;
                    }
                }
            }
        }
        __syncthreads();                           // every warp is done with K
        if (next < a.n_items) load_kv(next, 0);    // next item's K lands during softmax and PV
        else cp_async_commit();

        // ---- row maxima (rows g and g+8 live in the 4 lanes of a quad); the exponentials are taken inside the PV loop ----
        float inv0 = 0.f, inv1 = 0.f, m0 = 0.f, m1 = 0.f;
        if (active) {
            float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
            for (int t = 0; t < NT; ++t) {
                if (8 * t + 8 > T) {               // uniform: only the last tile(s) hold keys >= T
                    const int j = 8 * t + 2 * q;
#pragma unroll
                    for (int e = 0; e < 4; ++e) s[t][e] = (j + (e & 1) < T) ? s[t][e] : -INFINITY;
                }
                mx0 = fmaxf(mx0, fmaxf(s[t][0], s[t][1]));
                mx1 = fmaxf(mx1, fmaxf(s[t][2], s[t][3]));
            }
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
            m0 = mx0 * kscale;
            m1 = mx1 * kscale;
        }
        cp_async_wait<1>();                        // V of this item landed (the newest group is the next item's K)
        __syncthreads();

        // ---- O = P V : exp2((s - max) * scale) becomes the A fragments straight from the score registers, tile by tile under
        //      the mma of the previous step; 1/sum is applied to O ----
        float o[8][4];
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
        if (active) {
            const uint32_t v_lane = (uint32_t)(((lane & 7) + 8 * ((lane >> 3) & 1)) * ROWB + (lane >> 4) * 16);
            float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
            for (int kt = 0; kt < NKT; ++kt) {
#pragma unroll
                for (int t = 2 * kt; t < 2 * kt + 2; ++t) {
                    if (t < NT) {
                        s[t][0] = ex2(fmaf(s[t][0], kscale, -m0)); s[t][1] = ex2(fmaf(s[t][1], kscale, -m0));
                        s[t][2] = ex2(fmaf(s[t][2], kscale, -m1)); s[t][3] = ex2(fmaf(s[t][3], kscale, -m1));
                        sum0 += s[t][0] + s[t][1];
                        sum1 += s[t][2] + s[t][3];
                    }
                }
                uint32_t ph[4], pl[4];
                split2(s[2 * kt][0], s[2 * kt][1], ph[0], pl[0]);
                split2(s[2 * kt][2], s[2 * kt][3], ph[1], pl[1]);
                split2(s[2 * kt + 1][0], s[2 * kt + 1][1], ph[2], pl[2]);
                split2(s[2 * kt + 1][2], s[2 * kt + 1][3], ph[3], pl[3]);
                const uint32_t off = v_lane + (uint32_t)(16 * kt * ROWB);
                uint32_t vh[4][4], vl[4][4];               // [pair of 8-wide output tiles][b0,b1 | b0,b1]
#pragma unroll
                for (int np = 0; np < 4; ++np) {
                    ldsm4t(v_hi_s + off + np * 32, vh[np][0], vh[np][1], vh[np][2], vh[np][3]);
                    if (SPLIT) ldsm4t(v_lo_s + off + np * 32, vl[np][0], vl[np][1], vl[np][2], vl[np][3]);
                }
                // term by term over the 8 independent accumulators (dependency distance 8)
                if (SPLIT) {
#pragma unroll
                    for (int nt = 0; nt < 8; ++nt) mma16816(o[nt], pl, vh[nt >> 1][2 * (nt & 1)], vh[nt >> 1][2 * (nt & 1) + 1]);
#pragma unroll
                    for (int nt = 0; nt < 8; ++nt) mma16816(o[nt], ph, vl[nt >> 1][2 * (nt & 1)], vl[nt >> 1][2 * (nt & 1) + 1]);
                }
#pragma unroll
                for (int nt = 0; nt < 8; ++nt) mma16816(o[nt], ph, vh[nt >> 1][2 * (nt & 1)], vh[nt >> 1][2 * (nt & 1) + 1]);
            }
            sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1); sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
            sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1); sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
            inv0 = 1.f / sum0;
            inv1 = 1.f / sum1;
        }
        __syncthreads();                           // every warp is done with V
        if (next < a.n_items) {
            load_kv(next, 1);                      // next item's V lands during its score phase
            load_q(next);
        } else {
            cp_async_commit();
        }

        // ---- store rows i0+g, i0+g+8: fp32 and/or bf16 split planes (the out-projection GEMM's A operand) ----
        const int dm = H * DK;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int r = i0 + g + 8 * half;
            if (r >= T) continue;
            const long eoff = (n * T + r) * dm + h * DK + 2 * q;
            const float inv = half ? inv1 : inv0;
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                o[nt][2 * half] *= inv;
                o[nt][2 * half + 1] *= inv;
            }
            if (a.out) {
                float* dst = a.out + eoff;
#pragma unroll
                for (int nt = 0; nt < 8; ++nt) *reinterpret_cast<float2*>(dst + 8 * nt) = make_float2(o[nt][2 * half], o[nt][2 * half + 1]);
            }
            for (int p = 0; p < a.n_planes; ++p) {         // plane p takes the bf16 rounding of what planes < p left over
                __nv_bfloat16* dst = a.planes + p * a.plane_stride + eoff;
#pragma unroll
                for (int nt = 0; nt < 8; ++nt) {
                    const __nv_bfloat162 hb = __floats2bfloat162_rn(o[nt][2 * half], o[nt][2 * half + 1]);
                    o[nt][2 * half] -= __bfloat162float(hb.x);
                    o[nt][2 * half + 1] -= __bfloat162float(hb.y);
                    *reinterpret_cast<__nv_bfloat162*>(dst + 8 * nt) = hb;
                }
            }
        }
    }
    cp_async_wait<0>();
}

}  // namespace

bool attention_tc_supported(int T, int dk) { return dk == DK && T >= 1 && T <= TMAX; }

// qkv / pe given as bf16 split planes (in_planes = 1 or 2); out (fp32, optional) and/or bf16 planes (pitch H*64, optional)
int attention_tc_launch(const void* qkv_planes, long qkv_plane_stride, int in_planes, const void* pe_planes, long pe_plane_stride,
                        int pe_rows, float* out, void* planes, long plane_stride, int n_planes, long N, int T, int H, int dk,
                        cudaStream_t st) {
    CSS_REQUIRE(attention_tc_supported(T, dk), "attention_tc: T=%d / dk=%d unsupported", T, dk);
    CSS_REQUIRE(in_planes == 1 || in_planes == 2, "attention_tc: %d input planes unsupported", in_planes);
    CSS_REQUIRE(N > 0 && H > 0 && qkv_planes && (out || (planes && n_planes > 0)), "attention_tc: bad arguments");
    if (attention_umma_supported(T, dk, in_planes, pe_planes != nullptr, pe_rows)) {
        css_prof_begin(1, 6.0 * (double)N * H * T * T * DK, st);
        const int rc = attention_umma_launch(qkv_planes, qkv_plane_stride, pe_planes, pe_plane_stride, pe_rows, out, planes, plane_stride,
                                             planes ? n_planes : 0, N, H, st);
        css_prof_end(st);
        return rc;
    }
    typedef void (*Kern)(const AttnArgs);
    // score tiles per query tile: 19 covers the executor's 150-frame chunks (and anything up to 152), 20 the rest
    const bool split = in_planes == 2, use_pe = pe_planes != nullptr, nt19 = T <= 152;
    static const Kern kerns[8] = {attention_tc_kernel<false, false, 20>, attention_tc_kernel<false, false, 19>,
                                  attention_tc_kernel<false, true, 20>,  attention_tc_kernel<false, true, 19>,
                                  attention_tc_kernel<true, false, 20>,  attention_tc_kernel<true, false, 19>,
                                  attention_tc_kernel<true, true, 20>,   attention_tc_kernel<true, true, 19>};
    static CssPerDevice attr;
    if (attr.first()) {
        for (int i = 0; i < 8; ++i) {
            cudaError_t e = cudaFuncSetAttribute(kerns[i], cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
            CSS_REQUIRE(e == cudaSuccess, "attention_tc: cannot reserve %d B of shared memory: %s", SMEM_BYTES, cudaGetErrorString(e));
        }
    }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long items = N * H;
    const int grid = (int)(items < sms ? items : sms);
    AttnArgs a = {static_cast<const __nv_bfloat16*>(qkv_planes), qkv_plane_stride, static_cast<const __nv_bfloat16*>(pe_planes),
                  pe_plane_stride, pe_rows, out, static_cast<__nv_bfloat16*>(planes), plane_stride, planes ? n_planes : 0, items, T, H};
    css_prof_begin(1, 6.0 * (double)N * H * T * T * DK, st);   // QK^T + Q.PE^T + PV (algorithmic)
    kerns[4 * split + 2 * use_pe + nt19]<<<grid, NWARP * 32, SMEM_BYTES, st>>>(a);
    css_prof_end(st);
    return css_check_launch("attention_tc");
}

// fp32 entry (same contract as css_relpos_attention): the operands are split into two bf16 planes in
// stream-ordered scratch memory first.  The network path hands planes over directly.
extern "C" int css_relpos_attention_tc(const float* qkv, const float* pe, int pe_rows, float* out, long N, int T, int H, int dk,
                                       cudaStream_t st) {
    CSS_REQUIRE(attention_tc_supported(T, dk) && N > 0 && H > 0, "css_relpos_attention_tc: T=%d / dk=%d unsupported", T, dk);
    const long M = N * T;
    const int ld = 3 * H * dk;
    const size_t qb = (size_t)2 * M * ld * 2, pb = pe ? (size_t)2 * pe_rows * dk * 2 : 0;
    void* scratch = nullptr;
    cudaError_t e = cudaMallocAsync(&scratch, qb + pb + 256, st);
    CSS_REQUIRE(e == cudaSuccess, "css_relpos_attention_tc: scratch allocation failed: %s", cudaGetErrorString(e));
    PlaneMat qP = {scratch, M * (long)ld, ld, 2};
    PlaneMat pP = {static_cast<char*>(scratch) + ((qb + 255) & ~(size_t)255), (long)pe_rows * dk, dk, 2};
    int rc = split_planes_launch(qkv, M, ld, ld, qP, st);
    if (!rc && pe) rc = split_planes_launch(pe, pe_rows, dk, dk, pP, st);
    if (!rc) rc = attention_tc_launch(qP.ptr, qP.plane_stride, 2, pe ? pP.ptr : nullptr, pP.plane_stride, pe_rows, out, nullptr, 0, 0,
                                      N, T, H, dk, st);
    cudaFreeAsync(scratch, st);
    return rc;
}
