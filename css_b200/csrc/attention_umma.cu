// Relative-position multi-head self-attention core on tcgen05 / TMEM / TMA (css/models/conformer.py:229-250:
// A = Q K^T, B[i][j] = q_i . pe[i-j], softmax((A + B) / sqrt(dk)), P V) for the executor's 150-frame chunks.
//
// One persistent CTA per SM loops over (chunk, head) items; 18 warps:
//   warp 0      TMA producer: Q, K, V tiles of an item (bf16 hi/lo planes written by the QKV GEMM's epilogue) with
//               cp.async.bulk.tensor, 128-byte swizzle; the 299 relative-position rows stay resident in shared memory
//   warp 1      MMA issuer (one lane): per 128-row query tile  S = Q K^T (N = 160) and R = Q PE^T (N = 288 / 176) into TMEM,
//               later O = P V with P read FROM TMEM (tcgen05.mma with the A operand in tensor memory) and V as an MN-major
//               shared-memory operand (the [key][dk] tile as TMA delivers it).  Every product is the three-term bf16 split
//               hi.hi + hi.lo + lo.hi accumulated in fp32, as in the GEMMs.
//   warps 2-17  softmax: TMEM lane = query row; the 150 keys of a row are split over four warps (same lane quadrant, 40 keys
//               each).  A thread fetches its 40 scores and the 71 relative terms that contain its skewed window
//               R[i][i - j + 149] with tcgen05.ld, aligns the window with a register barrel shifter (shift = lane id: the
//               skew inside a warp), forms exp2((s + r - max) / 8 log2 e), exchanges (max, sum) with the three warps that
//               share its rows through shared memory, and writes P as packed bf16 hi / lo back into the score columns with
//               tcgen05.st.  After P V the same warps scale O by 1 / rowsum and store the output planes.
// A 150-row chunk is two query tiles (rows 0-127, then rows 128-149 in a second 128-row tile); TMEM holds
// S|P (160 columns) + R (288) + O (64) = 512 columns, so O of a tile drains while the next tile's S / R are computed.
#include "common.cuh"
#include "conformer.h"
#include "gemm.h"
#include "tma_map.h"
#include <cuda_bf16.h>
#include <cstdlib>

namespace {

constexpr int T_ = 150;                 // frames per chunk (the only length this kernel serves)
constexpr int DK = 64;
constexpr int KROWS = 160;              // key / value / query rows staged per item (150 + zero or next-chunk rows, masked)
constexpr int PE_ROWS = 304;            // relative positions c = i - j + 149 in [0, 298]
constexpr int NSPLIT = 4;               // warps per query row
constexpr int JW = 40;                  // keys per softmax thread
constexpr int NSOFT = 4 * NSPLIT;       // softmax warps
constexpr int THREADS = (2 + NSOFT) * 32;
constexpr int PLANE_T = KROWS * 128;    // one bf16 plane of Q, K or V: 160 rows x 128 B
constexpr int PLANE_PE = PE_ROWS * 128;
constexpr int OFF_PE = 0;
constexpr int OFF_Q = OFF_PE + 2 * PLANE_PE;
constexpr int OFF_K = OFF_Q + 2 * PLANE_T;
constexpr int OFF_V = OFF_K + 2 * PLANE_T;
constexpr int OFF_BAR = OFF_V + 2 * PLANE_T;
constexpr int OFF_XCH = OFF_BAR + 256;
constexpr int SMEM_BYTES = OFF_XCH + NSPLIT * 128 * 8 + 1024;
// TMEM columns
// O = P_hi [V_hi | V_lo] (one N = 128 MMA per k-step, the two V planes side by side along N) + P_lo V_hi (N = 64 into the first
// half): 128 columns that overlay the tail of R, which is dead once the softmax has read it.  (An N = 64 MMA costs the same
// ~70 cycles as an N = 128 one, so two MMAs per k-step instead of three N = 64 ones.)
constexpr int TM_S = 0, TM_PLO = 80, TM_R = 160, TM_O = 336;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] . B[smem]
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
// D[tmem] (+)= A[tmem] . B[smem]
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}\n" ::"r"(d), "r"(a), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
// K-major operand, rows of 64 bf16 = 128 B, SWIZZLE_128B: 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t desc_k(uint32_t saddr) {
    uint64_t d = (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// MN-major operand (V: [key][dk], dk contiguous, 64 bf16 = one 128-byte swizzle atom row): key groups of 8 are 1024 B apart
// (stride byte offset); 64-wide groups along N are `lbo_bytes` apart (leading byte offset: the distance between the hi and the lo
// plane when both are read as one N = 128 operand)
__device__ __forceinline__ uint64_t desc_mn(uint32_t saddr, uint32_t lbo_bytes = 1024) {
    uint64_t d = (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)(lbo_bytes >> 4) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// instruction descriptor: D = f32, A = B = bf16, M = 128, N; b_mn: B operand MN-major
__device__ __forceinline__ constexpr uint32_t idesc_of(int N, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

// ---- tcgen05.ld / st, 32 lanes x N consecutive 32-bit columns (thread t <-> TMEM lane base + t) ----
#define R4(a, i) "=r"(a[i]), "=r"(a[i + 1]), "=r"(a[i + 2]), "=r"(a[i + 3])
#define R8(a, i) R4(a, i), R4(a, i + 4)
#define R16(a, i) R8(a, i), R8(a, i + 8)
#define R32(a, i) R16(a, i), R16(a, i + 16)
#define I4(a, i) "r"(a[i]), "r"(a[i + 1]), "r"(a[i + 2]), "r"(a[i + 3])
#define I8(a, i) I4(a, i), I4(a, i + 4)
#define I16(a, i) I8(a, i), I8(a, i + 8)
__device__ __forceinline__ void tld8(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : R8(r, 0) : "r"(taddr));
}
__device__ __forceinline__ void tld32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : R32(r, 0) : "r"(taddr));
}
__device__ __forceinline__ void tld64(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,"
        "%32,%33,%34,%35,%36,%37,%38,%39,%40,%41,%42,%43,%44,%45,%46,%47,%48,%49,%50,%51,%52,%53,%54,%55,%56,%57,%58,%59,%60,%61,%62,%63}, [%64];"
        : R32(r, 0), R32(r, 32) : "r"(taddr));
}
__device__ __forceinline__ void tld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tst1(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr), "r"(r[0]) : "memory");
}
__device__ __forceinline__ void tst2(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1,%2};" ::"r"(taddr), "r"(r[0]), "r"(r[1]) : "memory");
}
__device__ __forceinline__ void tst4(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(taddr), I4(r, 0) : "memory");
}
__device__ __forceinline__ void tst8(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), I8(r, 0) : "memory");
}
__device__ __forceinline__ void tst16(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(taddr), I16(r, 0) : "memory");
}
__device__ __forceinline__ void tst_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// `ncols` (compile time) consecutive columns from r[]
template <int NC>
__device__ __forceinline__ void tst_n(uint32_t taddr, const uint32_t* r) {
    constexpr int O8 = NC & 16, O4 = NC & 24, O2 = NC & 28, O1 = NC & 30;
    if constexpr ((NC & 16) != 0) tst16(taddr, r);
    if constexpr ((NC & 8) != 0) tst8(taddr + O8, r + O8);
    if constexpr ((NC & 4) != 0) tst4(taddr + O4, r + O4);
    if constexpr ((NC & 2) != 0) tst2(taddr + O2, r + O2);
    if constexpr ((NC & 1) != 0) tst1(taddr + O1, r + O1);
    static_assert(NC < 32, "tst_n: at most 31 columns");
}
// register select that the compiler cannot turn into an indexed local-memory access
__device__ __forceinline__ uint32_t selp(uint32_t a, uint32_t b, uint32_t pred) {
    uint32_t d;
    asm("{\n.reg .pred p;\nsetp.ne.b32 p, %3, 0;\nselp.b32 %0, %1, %2, p;\n}" : "=r"(d) : "r"(a), "r"(b), "r"(pred));
    return d;
}

__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// one stage of the barrel shifter: r[k] <- r[k + SH] when bit log2(SH) of the lane id is set; JW + SH - 1 elements are still
// needed below this stage.  Fixed trip counts: every index is a compile-time constant and r[] stays in registers.
template <int SH>
__device__ __forceinline__ void shift_stage(uint32_t* r, int lane) {
    const uint32_t on = (uint32_t)lane & (uint32_t)SH;
#pragma unroll
    for (int k = 0; k < JW + SH - 1; ++k) r[k] = selp(r[k + SH], r[k], on);
}

struct UArgs {
    float* out;                 // (M, H*64) fp32 or null
    __nv_bfloat16* planes;      // [n_planes][M][H*64] or null
    long plane_stride;
    int n_planes;
    long n_items;
    int H;
    int pe_row0;                // table row of relative position c = 0 (i - j = -149)
    long long* dbg;             // optional: clock64 stamps of CTA 0 (css_attention_umma_debug), 16 per item for 8 items
};
#define DBG(slot) do { if (a.dbg && blockIdx.x == 0 && n < 8) a.dbg[n * 16 + (slot)] = clock64(); } while (0)

// Softmax of one query tile for the thread's row and key group G (keys [J1 - 40, J1), J1 = 150 - 40 G, clipped at 0).
//   window element m <-> key j = J1 - 1 - m <-> relative column c = cb + lane + m, cb = 32 q + 40 G
template <int G>
__device__ __forceinline__ void softmax_tile(uint32_t tmem, int q, int lane, float* xch, int row, float& inv_sum) {
    constexpr int J1 = T_ - JW * G;
    constexpr int JLO = J1 - JW < 0 ? 0 : J1 - JW;
    constexpr int NV = J1 - JLO;                       // valid keys: 40, or 30 in the last group
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const float kscale = 0.125f * 1.4426950408889634f;  // 1 / sqrt(64) * log2(e)
    float v[JW];
    {
        // relative terms: columns [cb, cb + 72) of R, window [lane, lane + 40) of them
        uint32_t r[72];
        const uint32_t cb = (uint32_t)(32 * q + JW * G);
        tld64(tmem + lane_base + TM_R + cb, r);
        tld8(tmem + lane_base + TM_R + cb + 64, r + 64);
        tld_wait();
        // barrel shifter: out[m] = r[m + lane]; every index below is a compile-time constant
        shift_stage<16>(r, lane);
        shift_stage<8>(r, lane);
        shift_stage<4>(r, lane);
        shift_stage<2>(r, lane);
        shift_stage<1>(r, lane);
        uint32_t s[JW];
        tld32(tmem + lane_base + TM_S + JLO, s);
        tld8(tmem + lane_base + TM_S + JLO + 32, s + 32);
        tld_wait();
#pragma unroll
        for (int m = 0; m < JW; ++m) {
            const int jj = NV - 1 - m;                 // s index of key j = J1 - 1 - m
            v[m] = (m < NV) ? __uint_as_float(s[jj < 0 ? 0 : jj]) + __uint_as_float(r[m]) : -INFINITY;
        }
    }
    float mx = -INFINITY;
#pragma unroll
    for (int m = 0; m < NV; ++m) mx = fmaxf(mx, v[m]);
    float sum = 0.f;
    const float mxs = mx * kscale;
#pragma unroll
    for (int m = 0; m < NV; ++m) {
        v[m] = ex2(fmaf(v[m], kscale, -mxs));
        sum += v[m];
    }
    // exchange (max, sum) with the three other key groups of this row
    xch[(G * 128 + row) * 2] = mx;
    xch[(G * 128 + row) * 2 + 1] = sum;
    asm volatile("bar.sync %0, %1;" ::"r"(1 + q), "n"(NSPLIT * 32) : "memory");     // the four warps that share these 32 rows
    float gm = mx, tot = 0.f;
#pragma unroll
    for (int g2 = 0; g2 < NSPLIT; ++g2) gm = fmaxf(gm, xch[(g2 * 128 + row) * 2]);
#pragma unroll
    for (int g2 = 0; g2 < NSPLIT; ++g2) tot += xch[(g2 * 128 + row) * 2 + 1] * ex2((xch[(g2 * 128 + row) * 2] - gm) * kscale);
    inv_sum = 1.f / tot;
    const float f = ex2((mx - gm) * kscale);
    // P (unnormalised, in [0, 1]) as packed bf16 hi / lo: column k of the plane holds keys 2k (low half), 2k + 1
    constexpr int NC = NV / 2;                         // 20 or 15 columns
    constexpr int NCW = (G == 0) ? NC + 5 : NC;        // group 0 also zeroes the columns of the masked keys 150..159
    uint32_t ph[NCW], pl[NCW];
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        const int j0 = JLO + 2 * k;                    // keys j0, j0 + 1 <-> m = J1 - 1 - j
        const float a = v[J1 - 1 - j0] * f, b = v[J1 - 2 - j0] * f;
        const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        const __nv_bfloat162 l = __floats2bfloat162_rn(a - __bfloat162float(h.x), b - __bfloat162float(h.y));
        ph[k] = *reinterpret_cast<const uint32_t*>(&h);
        pl[k] = *reinterpret_cast<const uint32_t*>(&l);
    }
#pragma unroll
    for (int k = NC; k < NCW; ++k) ph[k] = pl[k] = 0u;
    tst_n<NCW>(tmem + lane_base + TM_S + JLO / 2, ph);
    tst_n<NCW>(tmem + lane_base + TM_PLO + JLO / 2, pl);
    tst_wait();
}

__global__ void __launch_bounds__(THREADS, 1)
attention_umma_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmPE, const UArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t sb = smem_u32(smem);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
    uint64_t* pe_full = bars + 0;
    uint64_t* qk_full = bars + 1;
    uint64_t* v_full = bars + 2;
    uint64_t* qk_free = bars + 3;
    uint64_t* v_free = bars + 4;
    uint64_t* sr_full = bars + 5;
    uint64_t* p_full = bars + 6;
    uint64_t* o_full = bars + 7;
    uint64_t* o_free = bars + 8;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);
    float* xch = reinterpret_cast<float*>(smem + OFF_XCH);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int H = a.H;

    if (warp == 1 && lane == 0) {
        mbar_init(pe_full, 1); mbar_init(qk_full, 1); mbar_init(v_full, 1); mbar_init(qk_free, 1); mbar_init(v_free, 1);
        mbar_init(sr_full, 1); mbar_init(p_full, NSOFT); mbar_init(o_full, 1); mbar_init(o_free, 8);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmQKV)) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmPE)) : "memory");
            mbar_expect_tx(pe_full, 2 * PLANE_PE);
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                tma_load_3d(sb + OFF_PE + p * PLANE_PE, &tmPE, pe_full, 0, a.pe_row0, p);
                tma_load_3d(sb + OFF_PE + p * PLANE_PE + 152 * 128, &tmPE, pe_full, 0, a.pe_row0 + 152, p);
            }
            uint32_t n = 0;
            for (long it = blockIdx.x; it < a.n_items; it += gridDim.x, ++n) {
                const long chunk = it / H;
                const int h = (int)(it - chunk * H);
                const int row0 = (int)(chunk * T_);
                if (n) mbar_wait(qk_free, (n - 1) & 1);
                mbar_expect_tx(qk_full, 4 * PLANE_T);
#pragma unroll
                for (int p = 0; p < 2; ++p) {
                    tma_load_3d(sb + OFF_Q + p * PLANE_T, &tmQKV, qk_full, h * DK, row0, p);
                    tma_load_3d(sb + OFF_K + p * PLANE_T, &tmQKV, qk_full, (H + h) * DK, row0, p);
                }
                if (n) mbar_wait(v_free, (n - 1) & 1);
                mbar_expect_tx(v_full, 2 * PLANE_T);
#pragma unroll
                for (int p = 0; p < 2; ++p) tma_load_3d(sb + OFF_V + p * PLANE_T, &tmQKV, v_full, (2 * H + h) * DK, row0, p);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            mbar_wait(pe_full, 0);
            uint32_t n = 0, tiles = 0;          // tiles: query tiles issued so far (two per item)
            for (long it = blockIdx.x; it < a.n_items; it += gridDim.x, ++n) {
                mbar_wait(qk_full, n & 1);
                tc_fence_after();
                DBG(0);
#pragma unroll 1
                for (int tile = 0; tile < 2; ++tile, ++tiles) {
                    const uint32_t qa = sb + OFF_Q + tile * 128 * 128;          // rows 128 tile ..
                    if (tile == 0) {
                        // S and R of tile 0; tile 1's were issued behind P V of tile 0 (below)
                        // ---- S = Q K^T, N = 160
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
#pragma unroll
                            for (int tm = 0; tm < 3; ++tm) {
                                const int pa = tm == 2, pb = tm == 1;
                                mma_ss(tmem + TM_S, desc_k(qa + pa * PLANE_T + ks * 32), desc_k(sb + OFF_K + pb * PLANE_T + ks * 32),
                                       idesc_of(160, false), (ks | tm) != 0);
                            }
                        // R overlays the O columns of the previous tile: its epilogue must have drained them
                        if (tiles) { mbar_wait(o_free, (tiles - 1) & 1); tc_fence_after(); }
                        // ---- R = Q PE^T, relative columns [0, 288) as two N = 144 halves
#pragma unroll
                        for (int half = 0; half < 2; ++half)
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks)
#pragma unroll
                                for (int tm = 0; tm < 3; ++tm) {
                                    const int pa = tm == 2, pb = tm == 1;
                                    mma_ss(tmem + TM_R + half * 144, desc_k(qa + pa * PLANE_T + ks * 32),
                                           desc_k(sb + OFF_PE + pb * PLANE_PE + half * 144 * 128 + ks * 32), idesc_of(144, false),
                                           (ks | tm) != 0);
                                }
                        tc_commit(sr_full);
                    }
                    // ---- O = P V once the softmax warps have written P (and finished reading R)
                    DBG(1 + 4 * tile);
                    mbar_wait(p_full, tiles & 1);
                    DBG(2 + 4 * tile);
                    if (tile == 0) mbar_wait(v_full, n & 1);
                    if (tiles) mbar_wait(o_free, (tiles - 1) & 1);
                    DBG(3 + 4 * tile);
                    tc_fence_after();
#pragma unroll
                    for (int ks = 0; ks < KROWS / 16; ++ks) {
                        // P_hi . [V_hi | V_lo]: the two V planes are PLANE_T bytes apart = the operand's leading byte offset
                        mma_ts(tmem + TM_O, tmem + TM_S + ks * 8, desc_mn(sb + OFF_V + ks * 16 * 128, PLANE_T), idesc_of(128, true), ks != 0);
                        mma_ts(tmem + TM_O, tmem + TM_PLO + ks * 8, desc_mn(sb + OFF_V + ks * 16 * 128), idesc_of(64, true), 1);
                    }
                    tc_commit(o_full);
                    if (tile == 0) {
                        // tile 1 (rows 128..149): R needs relative columns [128, 299) -> N = 176 from PE row 128; S behind P V
                        const uint32_t qb = sb + OFF_Q + 128 * 128;
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
#pragma unroll
                            for (int tm = 0; tm < 3; ++tm) {
                                const int pa = tm == 2, pb = tm == 1;
                                mma_ss(tmem + TM_R, desc_k(qb + pa * PLANE_T + ks * 32),
                                       desc_k(sb + OFF_PE + pb * PLANE_PE + 128 * 128 + ks * 32), idesc_of(176, false), (ks | tm) != 0);
                            }
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
#pragma unroll
                            for (int tm = 0; tm < 3; ++tm) {
                                const int pa = tm == 2, pb = tm == 1;
                                mma_ss(tmem + TM_S, desc_k(qb + pa * PLANE_T + ks * 32), desc_k(sb + OFF_K + pb * PLANE_T + ks * 32),
                                       idesc_of(160, false), (ks | tm) != 0);
                            }
                        tc_commit(sr_full);
                        tc_commit(qk_free);              // Q and K of this item are no longer read
                    } else {
                        tc_commit(v_free);
                    }
                }
            }
        }
    } else {
        // ===== softmax / epilogue warps =====
        const int q = warp & 3;                   // TMEM lane quadrant this warp may access
        const int G = (warp - 2) >> 2;            // key group
        const int row = q * 32 + lane;            // row inside the query tile = TMEM lane
        const uint32_t lane_base = (uint32_t)(q * 32) << 16;
        uint32_t tiles = 0, n = 0;
        for (long it = blockIdx.x; it < a.n_items; it += gridDim.x, ++n) {
            const long chunk = it / H;
            const int h = (int)(it - chunk * H);
#pragma unroll 1
            for (int tile = 0; tile < 2; ++tile, ++tiles) {
                // tile 1 holds rows 128..149 only: the warps of lane quadrants 1-3 own no valid row there (their P rows feed
                // output rows that are never stored) and just keep the barrier counts
                // (they still wait for the phase's barrier before arriving: an early second arrival would be counted in the
                // phase of tile 0)
                const bool live = tile == 0 || q == 0;
                float inv_sum = 0.f;
                mbar_wait(sr_full, tiles & 1);
                if (live) {
                    tc_fence_after();
                    if (threadIdx.x == 128) DBG(9 + 3 * tile);
                    if (G == 0) softmax_tile<0>(tmem, q, lane, xch, row, inv_sum);
                    else if (G == 1) softmax_tile<1>(tmem, q, lane, xch, row, inv_sum);
                    else if (G == 2) softmax_tile<2>(tmem, q, lane, xch, row, inv_sum);
                    else softmax_tile<3>(tmem, q, lane, xch, row, inv_sum);
                    tc_fence_before();
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(p_full);
                if (threadIdx.x == 128) DBG(10 + 3 * tile);
                if (G >= 2) continue;
                if (!live) {
                    mbar_wait(o_full, tiles & 1);
                    if (lane == 0) mbar_arrive(o_free);
                    continue;
                }
                // ---- epilogue: key groups 0 and 1 take 32 output columns each
                mbar_wait(o_full, tiles & 1);
                tc_fence_after();
                if (threadIdx.x == 128) DBG(11 + 3 * tile);
                uint32_t o[32], o2[32];
                tld32(tmem + lane_base + TM_O + 32 * G, o);
                tld32(tmem + lane_base + TM_O + 64 + 32 * G, o2);
                tld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(o_free);
                const int t = tile * 128 + row;
                if (t < T_) {
                    const long eoff = (chunk * T_ + t) * (long)(H * DK) + h * DK + 32 * G;
                    float f[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) f[i] = (__uint_as_float(o[i]) + __uint_as_float(o2[i])) * inv_sum;
                    if (a.out) {
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            *reinterpret_cast<float4*>(a.out + eoff + 4 * i) = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
                    }
                    for (int p = 0; p < a.n_planes; ++p) {
                        uint32_t w[16];
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const __nv_bfloat162 hb = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
                            f[2 * i] -= __bfloat162float(hb.x);
                            f[2 * i + 1] -= __bfloat162float(hb.y);
                            w[i] = *reinterpret_cast<const uint32_t*>(&hb);
                        }
                        uint4* dst = reinterpret_cast<uint4*>(a.planes + p * a.plane_stride + eoff);
#pragma unroll
                        for (int i = 0; i < 4; ++i) dst[i] = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
    }
}

}  // namespace

static long long* g_dbg = nullptr;
extern "C" void css_attention_umma_debug(long long* buf) { g_dbg = buf; }

bool attention_umma_supported(int T, int dk, int in_planes, bool has_pe, int pe_rows) {
    static const bool off = getenv("CSS_B200_ATT_UMMA") && getenv("CSS_B200_ATT_UMMA")[0] == '0';
    return !off && T == T_ && dk == DK && in_planes == 2 && has_pe && pe_rows / 2 >= T_ && pe_rows / 2 + 155 <= pe_rows;
}

int attention_umma_launch(const void* qkv_planes, long qkv_plane_stride, const void* pe_planes, long pe_plane_stride, int pe_rows,
                          float* out, void* planes, long plane_stride, int n_planes, long N, int H, cudaStream_t st) {
    const long M = N * T_;
    const int ld = 3 * H * DK;
    CUtensorMap tq, tp;
    int rc;
    if ((rc = css_make_plane_map(&tq, qkv_planes, M, ld, ld, qkv_plane_stride, 2, KROWS, 64))) return rc;
    if ((rc = css_make_plane_map(&tp, pe_planes, pe_rows, DK, DK, pe_plane_stride, 2, 152, 64))) return rc;
    static CssPerDevice attr;
    if (attr.first()) {
        cudaError_t e = cudaFuncSetAttribute(attention_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        CSS_REQUIRE(e == cudaSuccess, "attention_umma: cannot reserve %d B of shared memory: %s", SMEM_BYTES, cudaGetErrorString(e));
    }
    const long items = N * H;
    const int grid = (int)(items < sm_count() ? items : sm_count());
    UArgs a = {out, static_cast<__nv_bfloat16*>(planes), plane_stride, planes ? n_planes : 0, items, H, pe_rows / 2 - (T_ - 1), g_dbg};
    attention_umma_kernel<<<grid, THREADS, SMEM_BYTES, st>>>(tq, tp, a);
    return css_check_launch("attention_umma");
}
