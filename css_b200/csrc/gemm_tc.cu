// tcgen05 / TMEM / TMA GEMM for the dense contractions of the mask networks
// (torch.nn.Linear call sites: css/models/conformer.py:79-81,124,205-209,314-320; css/models/blstm.py:100,110,139).
//
//   C[m][n] = epilogue( sum_k A[m][k] * W[n][k] )          A: (M,K) K-major,  W: (N,K) K-major
//
// Operands live in HBM as P "planes" of bf16 (split-float: x = p0 + p1 (+ p2), each plane the bf16 rounding of
// the remaining residue), written by the producing kernel (LayerNorm, attention, the previous GEMM's epilogue).
//   P = 1 : plain bf16                       1 MMA  per k-step
//   P = 2 : bf16x3  (p0q0 + p0q1 + p1q0)      3 MMAs per k-step, ~2^-17 relative product error
//   P = 3 : bf16x6  (all terms of order < 3)  6 MMAs per k-step, fp32-class accuracy (fp32 accumulation in TMEM)
//
// Structure (one persistent CTA per SM, 384 threads):
//   warp 0  TMA producer: cp.async.bulk.tensor.3d (k, row, plane) into a STAGES-deep smem ring, 128B swizzle
//   warp 1  MMA issuer : one elected lane issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=BN, K=16),
//                         accumulators in TMEM, double buffered (2 x BN columns) so the epilogue overlaps the
//                         next tile's main loop; tcgen05.commit releases smem slots / signals the epilogue
//   warp 2  TMEM allocator
//   warps 4-11 epilogue: tcgen05.ld (32 lanes x 32 columns per warp) -> swizzled smem transpose -> bias / ReLU /
//                         clamp / alpha / residual -> coalesced fp32 rows and/or bf16 planes for the next GEMM
#include "common.cuh"
#include "gemm.h"
#include <cuda.h>
#include <cstdlib>
#include <cuda_bf16.h>

namespace {

constexpr int BM = 128;
constexpr int KPAD = 64;          // K granularity of the plane buffers (both BK choices divide it)
constexpr int UMMA_K = 16;
constexpr int EPI_STAGING = 8 * 4096;      // 8 epilogue warps x (32 rows x 32 fp32)
constexpr int SMEM_BUDGET = 227 * 1024 - EPI_STAGING - 1024 - 256;

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
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// the epilogue's staging tile is addressed in the shared window explicitly: through the (re-aligned) generic pointer the
// compiler emitted generic LD.E/ST.E, whose long-scoreboard latency sat on the critical path of every chunk
__device__ __forceinline__ void sts128(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread t of the warp receives lane (base_lane + t)
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major operand tile with rows of BK bf16 (128 B -> SWIZZLE_128B, 64 B -> SWIZZLE_64B): 8-row groups are
// 8*row bytes apart (SBO), LBO unused (=1).
template <int BK>
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    static_assert(BK == 64 || BK == 32, "BK must be 64 (128B swizzle) or 32 (64B swizzle)");
    uint64_t d = (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;                              // leading byte offset (ignored for swizzled K-major)
    d |= (uint64_t)((8 * BK * 2) >> 4) << 32;            // stride byte offset
    d |= (uint64_t)1 << 46;                              // descriptor version (Blackwell)
    d |= (uint64_t)(BK == 64 ? 2 : 4) << 61;             // SWIZZLE_128B / SWIZZLE_64B
    return d;
}

// MN-major operand tile: 64-element (128 B) groups along M/N, each group = BK rows (k) of 128 B in the 128-byte swizzle;
// 8-row k groups are 1024 B apart (stride byte offset), the 64-wide groups BK * 128 B apart (leading byte offset).
template <int BK>
__device__ __forceinline__ uint64_t make_desc_mn(uint32_t saddr) {
    uint64_t d = (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((BK * 128) >> 4) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

struct TcEpilogue {
    const float* bias;
    int act;
    float alpha;
    const float* residual;
    int ldr;
    float clamp_max;
    float* c;                 // fp32 output (may be null)
    int ldc;
    __nv_bfloat16* planes;    // bf16 plane output (may be null)
    long plane_stride;        // elements between planes
    int ldp;
    int n_planes;
    // training extensions
    const __nv_bfloat16* mask;   // forward activation (plane 0): output zeroed where mask <= 0 (ReLU backward)
    int ldmask;
    int ksplit;                  // > 1: the K range is split over `ksplit` work items that accumulate into c atomically
    // EM 4: planes = split planes of LayerNorm(c row; ln_g, ln_b)
    const float* ln_g;
    const float* ln_b;
    int ln_to_c;                 // c = LayerNorm(result) and planes = split(c) or split(LayerNorm(c; ln2_g, ln2_b))
    const float* ln2_g;
    const float* ln2_b;
};

template <int P> struct Terms;
template <> struct Terms<1> { static constexpr int N = 1; };
template <> struct Terms<2> { static constexpr int N = 3; };
template <> struct Terms<3> { static constexpr int N = 6; };
// product terms in decreasing order of magnitude: (plane of A, plane of W)
__device__ __forceinline__ constexpr int term_a(int t) { return t == 2 ? 1 : (t == 4 ? 2 : (t == 5 ? 1 : 0)); }
__device__ __forceinline__ constexpr int term_b(int t) { return t == 1 ? 1 : (t == 3 ? 2 : (t == 5 ? 1 : 0)); }

template <int P, int BN, int BK>
struct Cfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = P * (A_BYTES + B_BYTES);
    static constexpr int STAGES_RAW = SMEM_BUDGET / STAGE_BYTES;
    static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
    static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/ + EPI_STAGING;
    static_assert(STAGES >= 2, "tile does not fit a 2-stage ring");
};

// The lane's four bias values of each 32-column chunk of a tile (lean epilogues), fetched by the caller BEFORE it waits for
// the accumulator: a load issued inside the chunk loop put one global-memory latency on the critical path of every chunk.
template <int BN, int EM>
__device__ __forceinline__ void epilogue_bias(const TcEpilogue& ep, int nb, int N, int chalf, int lane, float4 (&bbv)[BN / 64]) {
#pragma unroll
    for (int cc = 0; cc < BN / 64; ++cc) {
        const int col = nb * BN + chalf * (BN / 2) + cc * 32 + (lane & 7) * 4;
        bbv[cc] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (EM != 0 && ep.bias && col < N)
            bbv[cc] = make_float4(__ldg(ep.bias + col), __ldg(ep.bias + col + 1), __ldg(ep.bias + col + 2), __ldg(ep.bias + col + 3));
    }
}

// Epilogue of one 128 x BN accumulator (TMEM columns tmem_acc ..) by one of the 8 epilogue warps: warp (q, chalf) drains
// TMEM lanes 32q.. of the column half `chalf`.  row_base: global row of TMEM lane 32q; nb: n-tile index.
template <int P, int BN, bool TRAIN, int EM>
__device__ __forceinline__ void epilogue_tile(const TcEpilogue& ep, uint32_t tmem_acc, uint32_t stg, long row_base, int nb, long M,
                                              int N, int q, int chalf, int lane, int ksplit, const float4 (&bbv)[BN / 64]) {
    const int cg = lane & 7, rsub = lane >> 3;
#pragma unroll 1
    for (int cc = 0; cc < BN / 64; ++cc) {
        const int c0 = chalf * (BN / 2) + cc * 32;
        const int col0 = nb * BN + c0;
        if (col0 >= N && (!ep.planes || col0 >= ep.ldp)) break;          // warp-uniform
        // EM 2: the residual rows of the chunk are requested before the accumulator is fetched and transposed, so their
        // global-memory latency runs underneath that work (the residual may alias C: all loads precede the chunk's stores)
        float4 res4[8];
        if constexpr (EM == 2) {
            if (ep.residual) {
                const float* rp = ep.residual + (row_base + rsub) * ep.ldr + col0 + cg * 4;
#pragma unroll
                for (int it = 0; it < 8; ++it)
                    res4[it] = (row_base + rsub + it * 4 < M) ? *reinterpret_cast<const float4*>(rp + (long)(it * 4) * ep.ldr)
                                                              : make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
#pragma unroll
                for (int it = 0; it < 8; ++it) res4[it] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        uint32_t r[32];
        tc_ld32(tmem_acc + ((uint32_t)(q * 32) << 16) + c0, r);
        // phase A: thread = row; 16-byte pieces XOR-swizzled by the row so both phases are conflict-free
#pragma unroll
        for (int j = 0; j < 8; ++j)
            sts128(stg + (lane * 8 + (j ^ (lane & 7))) * 16, __uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                   __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
        __syncwarp();
        // phase B: 8 lanes cover one 128-byte row segment -> coalesced global traffic
        const int col = col0 + cg * 4;
        if constexpr (EM == 1 || EM == 2) {
            float4 xs[8];
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int rr = it * 4 + rsub;
                xs[it] = lds128(stg + (rr * 8 + (cg ^ (rr & 7))) * 16);
            }
            float4 bb = bbv[0];
#pragma unroll
            for (int i = 1; i < BN / 64; ++i) if (cc == i) bb = bbv[i];      // static indices: bbv stays in registers
            const long row0 = row_base + rsub;
            if constexpr (EM == 1) {
                const float floor_v = ep.act == 1 ? 0.f : -INFINITY;
                __nv_bfloat16* pp = ep.planes + row0 * ep.ldp + col;
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    if (row0 + it * 4 >= M) continue;
                    float v0 = fmaxf(xs[it].x + bb.x, floor_v), v1 = fmaxf(xs[it].y + bb.y, floor_v);
                    float v2 = fmaxf(xs[it].z + bb.z, floor_v), v3 = fmaxf(xs[it].w + bb.w, floor_v);
                    __nv_bfloat16* dst = pp + (long)(it * 4) * ep.ldp;
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        const __nv_bfloat162 h01 = __floats2bfloat162_rn(v0, v1);
                        const __nv_bfloat162 h23 = __floats2bfloat162_rn(v2, v3);
                        uint2 w;
                        w.x = *reinterpret_cast<const uint32_t*>(&h01);
                        w.y = *reinterpret_cast<const uint32_t*>(&h23);
                        *reinterpret_cast<uint2*>(dst + p * ep.plane_stride) = w;
                        if (p + 1 < P) {
                            v0 -= __uint_as_float(w.x << 16); v1 -= __uint_as_float(w.x & 0xffff0000u);
                            v2 -= __uint_as_float(w.y << 16); v3 -= __uint_as_float(w.y & 0xffff0000u);
                        }
                    }
                }
            } else {
                float* cp = ep.c + row0 * ep.ldc + col;
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    if (row0 + it * 4 >= M) continue;
                    float4 o;
                    o.x = fmaf(xs[it].x + bb.x, ep.alpha, res4[it].x);
                    o.y = fmaf(xs[it].y + bb.y, ep.alpha, res4[it].y);
                    o.z = fmaf(xs[it].z + bb.z, ep.alpha, res4[it].z);
                    o.w = fmaf(xs[it].w + bb.w, ep.alpha, res4[it].w);
                    *reinterpret_cast<float4*>(cp + (long)(it * 4) * ep.ldc) = o;
                }
            }
            __syncwarp();
            continue;
        }
        float b4[4] = {0.f, 0.f, 0.f, 0.f};
        if (ep.bias) {
#pragma unroll
            for (int i = 0; i < 4; ++i) if (col + i < N) b4[i] = __ldg(ep.bias + col + i);
        }
        const bool vec_ok = col + 4 <= N;
        // the residual may alias C (x <- x + ...): fetch all of this chunk's residual rows before any store,
        // otherwise every load would have to wait for the previous store
        if (ep.residual) {
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const long row = row_base + it * 4 + rsub;
                res4[it] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (row < M) {
                    const float* rp = ep.residual + row * ep.ldr + col;
                    if (vec_ok && (ep.ldr & 3) == 0) {
                        res4[it] = *reinterpret_cast<const float4*>(rp);
                    } else {
                        if (col < N) res4[it].x = rp[0];
                        if (col + 1 < N) res4[it].y = rp[1];
                        if (col + 2 < N) res4[it].z = rp[2];
                        if (col + 3 < N) res4[it].w = rp[3];
                    }
                }
            }
        }
        // all shared-memory reads of the chunk before the first global store: the compiler cannot prove that the
        // (generic) staging pointer and the outputs do not alias, so a load inside the store loop would serialise it
        float4 xs[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + rsub;
            xs[it] = lds128(stg + (rr * 8 + (cg ^ (rr & 7))) * 16);
        }
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + rsub;
            const long row = row_base + rr;
            const float4 x4 = xs[it];
            if (row >= M) continue;
            float v[4] = {x4.x + b4[0], x4.y + b4[1], x4.z + b4[2], x4.w + b4[3]};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (ep.act == 1) v[i] = fmaxf(v[i], 0.f);
                if (ep.clamp_max > 0.f) v[i] = fminf(v[i], ep.clamp_max);
                v[i] *= ep.alpha;
            }
            if (ep.residual) { v[0] += res4[it].x; v[1] += res4[it].y; v[2] += res4[it].z; v[3] += res4[it].w; }
            if (TRAIN && ep.mask) {
                const __nv_bfloat16* mp = ep.mask + row * ep.ldmask + col;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (col + i < N && !(__bfloat162float(mp[i]) > 0.f)) v[i] = 0.f;
            }
            if (TRAIN && ksplit > 1) {
                float* cp = ep.c + row * ep.ldc + col;
#pragma unroll
                for (int i = 0; i < 4; ++i) if (col + i < N) atomicAdd(cp + i, v[i]);
                continue;
            }
            if (ep.c) {
                float* cp = ep.c + row * ep.ldc + col;
                if (vec_ok && (ep.ldc & 3) == 0) {
                    *reinterpret_cast<float4*>(cp) = make_float4(v[0], v[1], v[2], v[3]);
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) if (col + i < N) cp[i] = v[i];
                }
            }
            if (ep.planes && col + 4 <= ep.ldp) {
                // split-float planes of the output (next GEMM's A operand); columns in [N, ldp) are zeroed
#pragma unroll
                for (int i = 0; i < 4; ++i) if (col + i >= N) v[i] = 0.f;
                for (int p = 0; p < ep.n_planes; ++p) {
                    const __nv_bfloat162 h01 = __floats2bfloat162_rn(v[0], v[1]);
                    const __nv_bfloat162 h23 = __floats2bfloat162_rn(v[2], v[3]);
                    v[0] -= __bfloat162float(h01.x); v[1] -= __bfloat162float(h01.y);
                    v[2] -= __bfloat162float(h23.x); v[3] -= __bfloat162float(h23.y);
                    uint2 w;
                    w.x = *reinterpret_cast<const uint32_t*>(&h01);
                    w.y = *reinterpret_cast<const uint32_t*>(&h23);
                    *reinterpret_cast<uint2*>(ep.planes + p * ep.plane_stride + row * ep.ldp + col) = w;
                }
            }
        }
        __syncwarp();
    }
}

// TRAIN = false is the inference instantiation: no split-K bookkeeping and no mask test in the epilogue (they cost the
// K = 256 GEMMs, whose tile time is set by the epilogue, 27 % when they were run-time branches)
// EM selects the epilogue: 0 generic; 1 bias (+ReLU) -> NPL = P bf16 planes only (FFN-in, QKV); 2 (acc + bias) * alpha
// + residual -> fp32 only (FFN-out, attention out-projection).  The lean forms need N % 32 == 0 and are chosen on the host;
// the K = 256 GEMMs are bound by epilogue instruction issue, so every instruction removed there is tile time.
// MN: 0 both operands K-major; 1 W MN-major (dX = dY W); 3 A and W MN-major (dW = dY^T X).  MN != 0 only with TRAIN.
template <int P, int BN, int BK, bool TRAIN, int EM, int MN = 0>
__global__ void __launch_bounds__(384, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, long M, int N, int K,
               TcEpilogue ep) {
    using C = Cfg<P, BN, BK>;
    constexpr int STAGES = C::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * C::STAGE_BYTES);
    uint64_t* full = bars;                    // [STAGES]
    uint64_t* empty = bars + STAGES;          // [STAGES]
    uint64_t* tfull = bars + 2 * STAGES;      // [2]
    uint64_t* tempty = bars + 2 * STAGES + 2; // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_tiles = (int)((M + BM - 1) / BM), n_tiles = (N + BN - 1) / BN;
    const int ksplit = TRAIN ? (ep.ksplit > 1 ? ep.ksplit : 1) : 1;
    const int tiles = m_tiles * n_tiles;
    const int total = tiles * ksplit;
    const int kblocks = K / BK;
    const int kper = (kblocks + ksplit - 1) / ksplit;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * BN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                const int tt = TRAIN ? t % tiles : t, sp = TRAIN ? t / tiles : 0;
                const int mb = tt / n_tiles, nb = tt - mb * n_tiles;
                const int kb0 = TRAIN ? sp * kper : 0, kb1 = TRAIN ? min(kblocks, kb0 + kper) : kblocks;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    mbar_expect_tx(&full[stage], C::STAGE_BYTES);
                    uint8_t* sa = smem + stage * C::STAGE_BYTES;
                    uint8_t* sb = sa + P * C::A_BYTES;
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        if constexpr ((MN & 2) != 0) {      // A^T planes (K, M): 64-column groups of BK rows
#pragma unroll
                            for (int g = 0; g < BM / 64; ++g)
                                tma_load_3d(sa + p * C::A_BYTES + g * BK * 128, &tmA, &full[stage], mb * BM + g * 64, kb * BK, p);
                        } else {
                            tma_load_3d(sa + p * C::A_BYTES, &tmA, &full[stage], kb * BK, mb * BM, p);
                        }
                        if constexpr ((MN & 1) != 0) {      // W^T planes (K, N)
#pragma unroll
                            for (int g = 0; g < BN / 64; ++g)
                                tma_load_3d(sb + p * C::B_BYTES + g * BK * 128, &tmB, &full[stage], nb * BN + g * 64, kb * BK, p);
                        } else {
                            tma_load_3d(sb + p * C::B_BYTES, &tmB, &full[stage], kb * BK, nb * BN, p);
                        }
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            // instruction descriptor: D=f32, A=B=bf16, K-major unless MN says otherwise (bits 15 / 16), N, M=128
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)((MN >> 1) & 1) << 15) | ((uint32_t)(MN & 1) << 16) |
                                   ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                mbar_wait(&tempty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                const int sp = TRAIN ? t / tiles : 0;
                const int kb0 = TRAIN ? sp * kper : 0, kb1 = TRAIN ? min(kblocks, kb0 + kper) : kblocks;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * C::STAGE_BYTES);
                    const uint32_t sb = sa + P * C::A_BYTES;
#pragma unroll
                    for (int ks = 0; ks < BK / UMMA_K; ++ks) {
#pragma unroll
                        for (int tm = 0; tm < Terms<P>::N; ++tm) {
                            const int pa = term_a(tm), pb = term_b(tm);
                            const uint64_t da = (MN & 2) ? make_desc_mn<BK>(sa + pa * C::A_BYTES + ks * UMMA_K * 128)
                                                         : make_desc<BK>(sa + pa * C::A_BYTES + ks * UMMA_K * 2);
                            const uint64_t db = (MN & 1) ? make_desc_mn<BK>(sb + pb * C::B_BYTES + ks * UMMA_K * 128)
                                                         : make_desc<BK>(sb + pb * C::B_BYTES + ks * UMMA_K * 2);
                            tc_mma(d_tmem, da, db, idesc, ((kb - kb0) | ks | tm) != 0);
                        }
                    }
                    tc_commit(&empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                tc_commit(&tfull[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue (8 warps): warp e drains TMEM lanes 32*(e%4).. of the column half e/4 =====
        const int e = warp - 4;
        const int q = e & 3;                  // TMEM lane quadrant == warp % 4
        const int chalf = e >> 2;
        const uint32_t stg = smem_u32(smem + STAGES * C::STAGE_BYTES + 256) + e * 4096;              // 4 KB per warp
        int acc = 0; uint32_t acc_phase = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x) {
            const int tt = TRAIN ? t % tiles : t;
            const int mb = tt / n_tiles, nb = tt - mb * n_tiles;
            float4 bbv[BN / 64];
            epilogue_bias<BN, EM>(ep, nb, N, chalf, lane, bbv);
            mbar_wait(&tfull[acc], acc_phase);
            tc_fence_after();
            epilogue_tile<P, BN, TRAIN, EM>(ep, tmem_base + acc * BN, stg, (long)mb * BM + q * 32, nb, M, N, q, chalf, lane, ksplit, bbv);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * BN));
    }
}

// ---- CTA-pair (cta_group::2) helpers ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local` (a shared::cta address of this CTA) in the CTA of rank `rank`
__device__ __forceinline__ uint32_t mapa_rank(uint32_t local, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load whose completion bytes are counted on a barrier that may live in the peer CTA of the pair
__device__ __forceinline__ void tma_load_3d_pair(void* dst, const CUtensorMap* map, uint32_t bar_cluster_addr, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// commit of the pair's MMAs: arrives on the barrier at the same offset in both CTAs
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc_mma_pair(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}

// ---- TMA-store epilogue (EM 3): bias (+ReLU) -> P bf16 planes, written by the bulk-copy engine ----
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(src), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void sts128u(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// One 128 x 256 accumulator -> planes.  The thread keeps the TMEM layout (lane = row, 32 consecutive columns): bias, ReLU and
// the bf16 split happen in registers, the 64-byte row pieces of each plane go to a 32 x 32 shared tile in the 64-byte
// swizzle of the output tensor map, and one lane hands the tile to the TMA store.  Compared with the transposing epilogue
// this leaves 4 STS.128 per plane per chunk on the LSU instead of 8 STS + 8 LDS + 16 STG: the K = 256 GEMMs were bound
// by the L1/LSU pipe (l1tex 72 % busy, stores backing up), not by the MMA.  Rows past M are clipped by the tensor map.
template <int P>
__device__ __forceinline__ void epilogue_tile_tma(const TcEpilogue& ep, const CUtensorMap* tmC, uint32_t tmem_acc, uint32_t stg,
                                                  long row_base, int nb, int q, int chalf, int lane) {
    constexpr int BN = 256;
    const float floor_v = ep.act == 1 ? 0.f : -INFINITY;
    const uint32_t sw = (uint32_t)((lane >> 1) & 3);           // SWIZZLE_64B: 16-byte piece j of row r sits at j ^ ((r >> 1) & 3)
#pragma unroll 1
    for (int cc = 0; cc < BN / 64; ++cc) {
        const int c0 = chalf * (BN / 2) + cc * 32;
        const int col0 = nb * BN + c0;
        // the chunk's 32 bias values, the same for every lane (L1-resident after the first tile)
        float4 bb[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
            bb[j] = ep.bias ? __ldg(reinterpret_cast<const float4*>(ep.bias + col0) + j) : make_float4(0.f, 0.f, 0.f, 0.f);
        uint32_t r[32];
        tc_ld32(tmem_acc + ((uint32_t)(q * 32) << 16) + c0, r);
        // the previous chunk's stores must have read the staging tile before it is overwritten (a second staging tile per
        // warp, so that a chunk need not wait for the previous chunk's store, was measured: no change)
        const uint32_t stb = stg;
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 4; ++j) {                           // 16-byte piece j: columns 8j .. 8j+7
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 b4 = bb[2 * j + (i >> 2)];
                const float bv = (i & 3) == 0 ? b4.x : ((i & 3) == 1 ? b4.y : ((i & 3) == 2 ? b4.z : b4.w));
                v[i] = fmaxf(__uint_as_float(r[8 * j + i]) + bv, floor_v);
            }
#pragma unroll
            for (int p = 0; p < P; ++p) {
                uint32_t w[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
                    w[i] = *reinterpret_cast<const uint32_t*>(&h);
                    if (p + 1 < P) {
                        v[2 * i] -= __uint_as_float(w[i] << 16);
                        v[2 * i + 1] -= __uint_as_float(w[i] & 0xffff0000u);
                    }
                }
                sts128u(stb + p * 2048 + lane * 64 + (((uint32_t)j ^ sw) << 4), w[0], w[1], w[2], w[3]);
            }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
#pragma unroll
            for (int p = 0; p < P; ++p) tma_store_3d(tmC, stb + p * 2048, col0, (int)row_base, p);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
}

// ---- EM 4: (acc + bias) * alpha + residual -> fp32 C, and the split planes of LayerNorm(C row) -> planes (N == 256) ----
// Pass 1 is the EM 2 epilogue plus per-row sums; the two warps that share a TMEM lane quadrant (column halves) swap their
// partial sums through shared memory across a 64-thread named barrier.  Pass 2 runs after the accumulator has been handed
// back to the MMA warp: every lane re-reads the C values it stored itself (L2), normalises and stores the planes.
struct LnRowStats { float mean[8], rstd[8]; };

template <int P>
__device__ __forceinline__ void epilogue_ln_pass1(const TcEpilogue& ep, uint32_t tmem_acc, uint32_t stg, uint32_t stg_partner,
                                                  long row_base, long M, int q, int chalf, int lane,
                                                  const float4 (&bbv)[4], LnRowStats& st) {
    constexpr int BN = 256;
    const int cg = lane & 7, rsub = lane >> 3;
    float s1[8], s2[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) s1[it] = s2[it] = 0.f;
    const long row0 = row_base + rsub;
#pragma unroll 1
    for (int cc = 0; cc < BN / 64; ++cc) {
        const int c0 = chalf * (BN / 2) + cc * 32;
        const int col = c0 + cg * 4;
        float4 res4[8];
        if (ep.residual) {
            const float* rp = ep.residual + row0 * ep.ldr + col;
#pragma unroll
            for (int it = 0; it < 8; ++it)
                res4[it] = (row0 + it * 4 < M) ? *reinterpret_cast<const float4*>(rp + (long)(it * 4) * ep.ldr) : make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
#pragma unroll
            for (int it = 0; it < 8; ++it) res4[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        uint32_t r[32];
        tc_ld32(tmem_acc + ((uint32_t)(q * 32) << 16) + c0, r);
#pragma unroll
        for (int j = 0; j < 8; ++j)
            sts128(stg + (lane * 8 + (j ^ (lane & 7))) * 16, __uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                   __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
        __syncwarp();
        float4 xs[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + rsub;
            xs[it] = lds128(stg + (rr * 8 + (cg ^ (rr & 7))) * 16);
        }
        float4 bb = bbv[0];
#pragma unroll
        for (int i = 1; i < 4; ++i) if (cc == i) bb = bbv[i];
        float* cp = ep.c + row0 * ep.ldc + col;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            float4 o;
            o.x = fmaf(xs[it].x + bb.x, ep.alpha, res4[it].x);
            o.y = fmaf(xs[it].y + bb.y, ep.alpha, res4[it].y);
            o.z = fmaf(xs[it].z + bb.z, ep.alpha, res4[it].z);
            o.w = fmaf(xs[it].w + bb.w, ep.alpha, res4[it].w);
            s1[it] += (o.x + o.y) + (o.z + o.w);
            s2[it] += (o.x * o.x + o.y * o.y) + (o.z * o.z + o.w * o.w);
            if (row0 + it * 4 < M) *reinterpret_cast<float4*>(cp + (long)(it * 4) * ep.ldc) = o;
        }
        __syncwarp();
    }
    // row sums over this warp's 128 columns: the 8 lanes of a row group hold the pieces
#pragma unroll
    for (int it = 0; it < 8; ++it) {
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            s1[it] += __shfl_xor_sync(0xffffffffu, s1[it], o);
            s2[it] += __shfl_xor_sync(0xffffffffu, s2[it], o);
        }
    }
    if (cg == 0) {
#pragma unroll
        for (int it = 0; it < 8; ++it)
            asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(stg + (it * 4 + rsub) * 8), "f"(s1[it]), "f"(s2[it]) : "memory");
    }
    asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        float a, b;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(a), "=f"(b) : "r"(stg_partner + (it * 4 + rsub) * 8) : "memory");
        const float mean = (s1[it] + a) * (1.f / BN);
        const float var = fmaxf((s2[it] + b) * (1.f / BN) - mean * mean, 0.f);
        st.mean[it] = mean;
        st.rstd[it] = rsqrtf(var + 1e-5f);
    }
    asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");      // both warps have read before the staging tiles are reused
}

// ---- ln_to_c form: C = LayerNorm(result; ln_g, ln_b) and planes = split(C) or split(LayerNorm(C; ln2_g, ln2_b)) ----
// The raw result is needed three times (row statistics, statistics of the normalised row, the stores) and never leaves the SM:
// pass 1 adds bias and residual in the row-major layout (coalesced residual reads through the staging tile, as above) and puts
// the values BACK into the accumulator's TMEM columns (tcgen05.st); the second norm's statistics are then thread-local sums in
// the TMEM layout (lane = row) and the last pass goes through the staging tile once more for coalesced stores.  L2 traffic per
// 128 x 256 tile: residual in, C and the planes out -- the unfused epilogue plus the planes -- instead of two more round trips of C.
__device__ __forceinline__ void tc_st32(uint32_t taddr, const uint32_t* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
          "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]),
          "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]),
          "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tc_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

template <int P>
__device__ __forceinline__ void epilogue_lnc(const TcEpilogue& ep, uint32_t tmem_acc, uint32_t stg, uint32_t stg_partner, long row_base,
                                             long M, int q, int chalf, int lane, const float4 (&bbv)[4]) {
    constexpr int BN = 256;
    const int cg = lane & 7, rsub = lane >> 3;
    const long row0 = row_base + rsub;
    const uint32_t tq = tmem_acc + ((uint32_t)(q * 32) << 16);
    float s1[8], s2[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) s1[it] = s2[it] = 0.f;
    // ---- pass 1: c = (acc + bias) * alpha + residual, back into TMEM; row sums ----
#pragma unroll 1
    for (int cc = 0; cc < BN / 64; ++cc) {
        const int c0 = chalf * (BN / 2) + cc * 32;
        const int col = c0 + cg * 4;
        float4 res4[8];
        if (ep.residual) {
            const float* rp = ep.residual + row0 * ep.ldr + col;
#pragma unroll
            for (int it = 0; it < 8; ++it)
                res4[it] = (row0 + it * 4 < M) ? *reinterpret_cast<const float4*>(rp + (long)(it * 4) * ep.ldr) : make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
#pragma unroll
            for (int it = 0; it < 8; ++it) res4[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        uint32_t r[32];
        tc_ld32(tq + c0, r);
#pragma unroll
        for (int j = 0; j < 8; ++j)
            sts128(stg + (lane * 8 + (j ^ (lane & 7))) * 16, __uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                   __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
        __syncwarp();
        float4 bb = bbv[0];
#pragma unroll
        for (int i = 1; i < 4; ++i) if (cc == i) bb = bbv[i];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + rsub;
            const uint32_t sa = stg + (rr * 8 + (cg ^ (rr & 7))) * 16;
            const float4 xs = lds128(sa);
            float4 o;
            o.x = fmaf(xs.x + bb.x, ep.alpha, res4[it].x);
            o.y = fmaf(xs.y + bb.y, ep.alpha, res4[it].y);
            o.z = fmaf(xs.z + bb.z, ep.alpha, res4[it].z);
            o.w = fmaf(xs.w + bb.w, ep.alpha, res4[it].w);
            s1[it] += (o.x + o.y) + (o.z + o.w);
            s2[it] += (o.x * o.x + o.y * o.y) + (o.z * o.z + o.w * o.w);
            sts128(sa, o.x, o.y, o.z, o.w);           // the same thread read this piece: no hazard inside the pass
        }
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 v = lds128(stg + (lane * 8 + (j ^ (lane & 7))) * 16);
            r[4 * j] = __float_as_uint(v.x); r[4 * j + 1] = __float_as_uint(v.y);
            r[4 * j + 2] = __float_as_uint(v.z); r[4 * j + 3] = __float_as_uint(v.w);
        }
        tc_st32(tq + c0, r);
        __syncwarp();
    }
    tc_st_wait();
#pragma unroll
    for (int it = 0; it < 8; ++it) {
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            s1[it] += __shfl_xor_sync(0xffffffffu, s1[it], o);
            s2[it] += __shfl_xor_sync(0xffffffffu, s2[it], o);
        }
    }
    if (cg == 0) {
#pragma unroll
        for (int it = 0; it < 8; ++it)
            asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(stg + (it * 4 + rsub) * 8), "f"(s1[it]), "f"(s2[it]) : "memory");
    }
    asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
    // statistics of the thread's OWN row (TMEM layout: lane = row) -- both warps of the pair compute the same numbers
    float mean1, rstd1;
    {
        float a0, b0, a1, b1;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(a0), "=f"(b0) : "r"(stg + lane * 8) : "memory");
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(a1), "=f"(b1) : "r"(stg_partner + lane * 8) : "memory");
        const float sa = chalf ? a1 + a0 : a0 + a1, sb = chalf ? b1 + b0 : b0 + b1;      // same order in both warps
        mean1 = sa * (1.f / BN);
        rstd1 = rsqrtf(fmaxf(sb * (1.f / BN) - mean1 * mean1, 0.f) + 1e-5f);
    }
    asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");      // both warps have read before the staging tiles are reused
    // ---- pass 2a: statistics of y = LayerNorm(c) for the second norm, thread-local in the TMEM layout ----
    float mean2 = 0.f, rstd2 = 1.f;
    const bool second = ep.ln2_g != nullptr;
    if (second) {
        float t1 = 0.f, t2 = 0.f;
#pragma unroll 1
        for (int cc = 0; cc < BN / 64; ++cc) {
            const int c0 = chalf * (BN / 2) + cc * 32;
            uint32_t r[32];
            tc_ld32(tq + c0, r);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 g4 = __ldg(reinterpret_cast<const float4*>(ep.ln_g + c0) + j);
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(ep.ln_b + c0) + j);
                const float y0 = (__uint_as_float(r[4 * j]) - mean1) * rstd1 * g4.x + b4.x;
                const float y1 = (__uint_as_float(r[4 * j + 1]) - mean1) * rstd1 * g4.y + b4.y;
                const float y2 = (__uint_as_float(r[4 * j + 2]) - mean1) * rstd1 * g4.z + b4.z;
                const float y3 = (__uint_as_float(r[4 * j + 3]) - mean1) * rstd1 * g4.w + b4.w;
                t1 += (y0 + y1) + (y2 + y3);
                t2 += (y0 * y0 + y1 * y1) + (y2 * y2 + y3 * y3);
            }
        }
        asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(stg + lane * 8), "f"(t1), "f"(t2) : "memory");
        asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
        float a1, b1;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(a1), "=f"(b1) : "r"(stg_partner + lane * 8) : "memory");
        const float sa = chalf ? a1 + t1 : t1 + a1, sb = chalf ? b1 + t2 : t2 + b1;
        mean2 = sa * (1.f / BN);
        rstd2 = rsqrtf(fmaxf(sb * (1.f / BN) - mean2 * mean2, 0.f) + 1e-5f);
        asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
    }
    // the four per-row numbers, from the TMEM layout to the row-major layout of the stores (8 rows per lane)
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(stg + lane * 16), "f"(mean1), "f"(rstd1), "f"(mean2), "f"(rstd2) : "memory");
    __syncwarp();
    float4 rs[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) rs[it] = lds128(stg + (it * 4 + rsub) * 16);
    __syncwarp();
    // ---- pass 2b: C = y, planes = split(y) or split(LayerNorm(y; ln2)) ----
#pragma unroll 1
    for (int cc = 0; cc < BN / 64; ++cc) {
        const int c0 = chalf * (BN / 2) + cc * 32;
        const int col = c0 + cg * 4;
        uint32_t r[32];
        tc_ld32(tq + c0, r);
#pragma unroll
        for (int j = 0; j < 8; ++j)
            sts128(stg + (lane * 8 + (j ^ (lane & 7))) * 16, __uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                   __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
        __syncwarp();
        const float4 g4 = __ldg(reinterpret_cast<const float4*>(ep.ln_g + col));
        const float4 b4 = __ldg(reinterpret_cast<const float4*>(ep.ln_b + col));
        float4 h4 = make_float4(1.f, 1.f, 1.f, 1.f), k4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (second) {
            h4 = __ldg(reinterpret_cast<const float4*>(ep.ln2_g + col));
            k4 = __ldg(reinterpret_cast<const float4*>(ep.ln2_b + col));
        }
        float* cp = ep.c + row0 * ep.ldc + col;
        __nv_bfloat16* pp = ep.planes + row0 * ep.ldp + col;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + rsub;
            const float4 x = lds128(stg + (rr * 8 + (cg ^ (rr & 7))) * 16);
            if (row0 + it * 4 >= M) continue;
            const float m = rs[it].x, rd = rs[it].y;
            float v0 = (x.x - m) * rd * g4.x + b4.x, v1 = (x.y - m) * rd * g4.y + b4.y;
            float v2 = (x.z - m) * rd * g4.z + b4.z, v3 = (x.w - m) * rd * g4.w + b4.w;
            *reinterpret_cast<float4*>(cp + (long)(it * 4) * ep.ldc) = make_float4(v0, v1, v2, v3);
            if (second) {
                const float m2 = rs[it].z, rd2 = rs[it].w;
                v0 = (v0 - m2) * rd2 * h4.x + k4.x; v1 = (v1 - m2) * rd2 * h4.y + k4.y;
                v2 = (v2 - m2) * rd2 * h4.z + k4.z; v3 = (v3 - m2) * rd2 * h4.w + k4.w;
            }
            __nv_bfloat16* dst = pp + (long)(it * 4) * ep.ldp;
#pragma unroll
            for (int p = 0; p < P; ++p) {
                const __nv_bfloat162 h01 = __floats2bfloat162_rn(v0, v1);
                const __nv_bfloat162 h23 = __floats2bfloat162_rn(v2, v3);
                uint2 w;
                w.x = *reinterpret_cast<const uint32_t*>(&h01);
                w.y = *reinterpret_cast<const uint32_t*>(&h23);
                *reinterpret_cast<uint2*>(dst + p * ep.plane_stride) = w;
                if (p + 1 < P) {
                    v0 -= __uint_as_float(w.x << 16); v1 -= __uint_as_float(w.x & 0xffff0000u);
                    v2 -= __uint_as_float(w.y << 16); v3 -= __uint_as_float(w.y & 0xffff0000u);
                }
            }
        }
        __syncwarp();
    }
}

template <int P>
__device__ __forceinline__ void epilogue_ln_pass2(const TcEpilogue& ep, const float* __restrict__ ln_g, const float* __restrict__ ln_b,
                                                  long row_base, long M, int chalf, int lane, const LnRowStats& st) {
    constexpr int BN = 256;
    const int cg = lane & 7, rsub = lane >> 3;
    const long row0 = row_base + rsub;
#pragma unroll 1
    for (int cc = 0; cc < BN / 64; ++cc) {
        const int col = chalf * (BN / 2) + cc * 32 + cg * 4;
        const float4 g4 = make_float4(__ldg(ln_g + col), __ldg(ln_g + col + 1), __ldg(ln_g + col + 2), __ldg(ln_g + col + 3));
        const float4 b4 = make_float4(__ldg(ln_b + col), __ldg(ln_b + col + 1), __ldg(ln_b + col + 2), __ldg(ln_b + col + 3));
        const float* cp = ep.c + row0 * ep.ldc + col;
        float4 x[8];
#pragma unroll
        for (int it = 0; it < 8; ++it)
            x[it] = (row0 + it * 4 < M) ? __ldcg(reinterpret_cast<const float4*>(cp + (long)(it * 4) * ep.ldc)) : make_float4(0.f, 0.f, 0.f, 0.f);
        __nv_bfloat16* pp = ep.planes + row0 * ep.ldp + col;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            if (row0 + it * 4 >= M) continue;
            const float m = st.mean[it], rs = st.rstd[it];
            float v0 = (x[it].x - m) * rs * g4.x + b4.x, v1 = (x[it].y - m) * rs * g4.y + b4.y;
            float v2 = (x[it].z - m) * rs * g4.z + b4.z, v3 = (x[it].w - m) * rs * g4.w + b4.w;
            __nv_bfloat16* dst = pp + (long)(it * 4) * ep.ldp;
#pragma unroll
            for (int p = 0; p < P; ++p) {
                const __nv_bfloat162 h01 = __floats2bfloat162_rn(v0, v1);
                const __nv_bfloat162 h23 = __floats2bfloat162_rn(v2, v3);
                uint2 w;
                w.x = *reinterpret_cast<const uint32_t*>(&h01);
                w.y = *reinterpret_cast<const uint32_t*>(&h23);
                *reinterpret_cast<uint2*>(dst + p * ep.plane_stride) = w;
                if (p + 1 < P) {
                    v0 -= __uint_as_float(w.x << 16); v1 -= __uint_as_float(w.x & 0xffff0000u);
                    v2 -= __uint_as_float(w.y << 16); v3 -= __uint_as_float(w.y & 0xffff0000u);
                }
            }
        }
    }
}

template <int P, int BK>
struct Cfg2 {
    static constexpr int BN = 256;
    static constexpr int A_BYTES = BM * BK * 2;           // this CTA's 128 rows of A
    static constexpr int B_BYTES = (BN / 2) * BK * 2;     // this CTA's half of the 256 W rows
    static constexpr int STAGE_BYTES = P * (A_BYTES + B_BYTES);
    static constexpr int STAGES_RAW = SMEM_BUDGET / STAGE_BYTES;
    static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
    static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/ + EPI_STAGING;
};

// CTA-pair form of the kernel above for the inference GEMMs with N % 256 == 0: two CTAs on the SMs of one TPC work on a
// 256 x 256 tile.  Each loads its own 128 rows of A and HALF of the W tile; the leader's tcgen05.mma.cta_group::2 (M = 256)
// reads both halves, so the W traffic into and out of shared memory per SM is halved -- shared-memory bandwidth (TMA fill
// + tensor-core operand reads + the epilogue's staging) is what bounds the one-CTA kernel, not the MMA rate.
// Each CTA's TMEM holds its own 128 rows of the accumulator and its own epilogue warps drain it.
template <int P, int BK, int EM>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(384, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmC, long M, int N, int K, TcEpilogue ep) {
    using C = Cfg2<P, BK>;
    constexpr int BN = 256;
    constexpr int STAGES = C::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    // the epilogue's staging tiles come first (1024-aligned: the TMA store of EM 3 reads them in the 64-byte swizzle)
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * C::STAGE_BYTES + EPI_STAGING);
    uint64_t* full = bars;                    // [STAGES]  used in the leader only (bytes of both CTAs)
    uint64_t* empty = bars + STAGES;          // [STAGES]  one per CTA, released by the pair's commit
    uint64_t* tfull = bars + 2 * STAGES;      // [2]       one per CTA
    uint64_t* tempty = bars + 2 * STAGES + 2; // [2]       used in the leader only: 16 epilogue warps of the pair
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    const int m_tiles = (int)((M + 2 * BM - 1) / (2 * BM)), n_tiles = N / BN;
    const int total = m_tiles * n_tiles;
    const int kblocks = K / BK;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 16); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * BN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer (both CTAs): own A rows, own half of W; bytes are counted on the leader's barrier =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int t = pair; t < total; t += n_pairs) {
                const int mb = t / n_tiles, nb = t - mb * n_tiles;
                for (int kb = 0; kb < kblocks; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    if (rank == 0) mbar_expect_tx(&full[stage], 2 * C::STAGE_BYTES);
                    const uint32_t fb = mapa_rank(smem_u32(&full[stage]), 0);
                    uint8_t* sa = smem + stage * C::STAGE_BYTES;
                    uint8_t* sb = sa + P * C::A_BYTES;
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        tma_load_3d_pair(sa + p * C::A_BYTES, &tmA, fb, kb * BK, mb * 2 * BM + (int)rank * BM, p);
                        tma_load_3d_pair(sb + p * C::B_BYTES, &tmB, fb, kb * BK, nb * BN + (int)rank * (BN / 2), p);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader only) =====
        if (lane == 0 && rank == 0) {
            // instruction descriptor: D=f32, A=B=bf16, both K-major, N = 256, M = 256 over the pair
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (int t = pair; t < total; t += n_pairs) {
                mbar_wait(&tempty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < kblocks; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * C::STAGE_BYTES);
                    const uint32_t sb = sa + P * C::A_BYTES;
#pragma unroll
                    for (int ks = 0; ks < BK / UMMA_K; ++ks) {
#pragma unroll
                        for (int tm = 0; tm < Terms<P>::N; ++tm) {
                            const int pa = term_a(tm), pb = term_b(tm);
                            const uint64_t da = make_desc<BK>(sa + pa * C::A_BYTES + ks * UMMA_K * 2);
                            const uint64_t db = make_desc<BK>(sb + pb * C::B_BYTES + ks * UMMA_K * 2);
                            tc_mma_pair(d_tmem, da, db, idesc, (kb | ks | tm) != 0);
                        }
                    }
                    tc_commit_pair(&empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                tc_commit_pair(&tfull[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue (8 warps per CTA): this CTA's 128 rows of the pair's tile =====
        const int e = warp - 4;
        const int q = e & 3;
        const int chalf = e >> 2;
        const uint32_t stg0 = smem_u32(smem + STAGES * C::STAGE_BYTES);
        const uint32_t stg = stg0 + e * 4096, stg_partner = stg0 + (e ^ 4) * 4096;    // partner: same rows, other column half
        if (EM == 3 && lane == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmC)) : "memory");
        int acc = 0; uint32_t acc_phase = 0;
        for (int t = pair; t < total; t += n_pairs) {
            const int mb = t / n_tiles, nb = t - mb * n_tiles;
            const long row_base = (long)mb * 2 * BM + (long)rank * BM + q * 32;
            if constexpr (EM == 3) {
                mbar_wait(&tfull[acc], acc_phase);
                tc_fence_after();
                epilogue_tile_tma<P>(ep, &tmC, tmem_base + acc * BN, stg, row_base, nb, q, chalf, lane);
            } else if constexpr (EM == 4) {
                float4 bbv[BN / 64];
                epilogue_bias<BN, EM>(ep, nb, N, chalf, lane, bbv);
                mbar_wait(&tfull[acc], acc_phase);
                tc_fence_after();
                if (ep.ln_to_c) {
                    // normalised C: the values stay in the accumulator's columns until the last pass has read them
                    epilogue_lnc<P>(ep, tmem_base + acc * BN, stg, stg_partner, row_base, M, q, chalf, lane, bbv);
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(mapa_rank(smem_u32(&tempty[acc]), 0));
                } else {
                    LnRowStats st;
                    epilogue_ln_pass1<P>(ep, tmem_base + acc * BN, stg, stg_partner, row_base, M, q, chalf, lane, bbv, st);
                    // the accumulator goes back to the MMA warp before the normalisation pass
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(mapa_rank(smem_u32(&tempty[acc]), 0));
                    epilogue_ln_pass2<P>(ep, ep.ln_g, ep.ln_b, row_base, M, chalf, lane, st);
                }
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
                continue;
            } else {
                float4 bbv[BN / 64];
                epilogue_bias<BN, EM>(ep, nb, N, chalf, lane, bbv);
                mbar_wait(&tfull[acc], acc_phase);
                tc_fence_after();
                epilogue_tile<P, BN, false, EM>(ep, tmem_base + acc * BN, stg, row_base, nb, M, N, q, chalf, lane, 1, bbv);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa_rank(smem_u32(&tempty[acc]), 0));
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
        if (EM == 3 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // stores done before exit
    }
    tc_fence_before();
    cluster_sync_all();          // the peer's shared memory and barriers stay valid until both CTAs are done
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * BN));
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
            qr == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// planes of a (rows, K) K-major bf16 matrix: element (p, r, k) at base[p*plane_stride + r*ld + k]
static int make_map(CUtensorMap* map, const void* base, long rows, int K, int ld, long plane_stride, int planes, int box_rows,
                    int BK) {
    EncodeTiledFn fn = encode_fn();
    CSS_REQUIRE(fn, "gemm_tc: cuTensorMapEncodeTiled is not available from the driver");
    CSS_REQUIRE(((uintptr_t)base & 15) == 0 && (ld * 2) % 16 == 0 && (plane_stride * 2) % 16 == 0,
                "gemm_tc: operand planes must be 16-byte aligned (ld=%d plane_stride=%ld)", ld, plane_stride);
    cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)rows, (cuuint64_t)planes};
    cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)plane_stride * 2};
    cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, BK == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CSS_REQUIRE(r == CUDA_SUCCESS, "gemm_tc: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return 0;
}

int css_make_f32_map3(CUtensorMap* map, const void* base, const unsigned long long dims[3], const unsigned long long strides[2],
                      const unsigned box[3]) {
    EncodeTiledFn fn = encode_fn();
    CSS_REQUIRE(fn, "cuTensorMapEncodeTiled is not available from the driver");
    CSS_REQUIRE(((uintptr_t)base & 15) == 0 && strides[0] % 16 == 0 && strides[1] % 16 == 0, "tensor map: base / strides must be 16-byte aligned");
    cuuint64_t d[3] = {dims[0], dims[1], dims[2]};
    cuuint64_t st[2] = {strides[0], strides[1]};
    cuuint32_t bx[3] = {box[0], box[1], box[2]};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(base), d, st, bx, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CSS_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (fp32, 3-D) failed (%d)", (int)r);
    return 0;
}

// shared with the tcgen05 attention kernel (attention_umma.cu)
int css_make_plane_map(CUtensorMap* map, const void* base, long rows, int K, int ld, long plane_stride, int planes, int box_rows,
                       int BK) {
    return make_map(map, base, rows, K, ld, plane_stride, planes, box_rows, BK);
}

int sm_count() {
    static int n_dev[CSS_MAX_DEVICES] = {};
    const int dev = css_device();
    int& n = n_dev[dev];
    if (!n) {
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

static int g_lean_epilogue = 1;   // 0: always the generic epilogue (A/B switch)
extern "C" void css_gemm_tc_lean_epilogue(int on) { g_lean_epilogue = on; }

static int g_two_cta = -1;        // CTA-pair kernel for N % 256 == 0 inference GEMMs; -1: read CSS_B200_GEMM_2CTA (default on)
extern "C" void css_gemm_tc_two_cta(int on) { g_two_cta = on ? 1 : 0; }
static bool two_cta_enabled() {
    if (g_two_cta < 0) {
        const char* e = getenv("CSS_B200_GEMM_2CTA");
        g_two_cta = (e && e[0] == '0') ? 0 : 1;
    }
    return g_two_cta != 0;
}

static int g_tma_epilogue = 1;    // planes-only epilogue of the CTA-pair kernel through the TMA store (A/B switch)
extern "C" void css_gemm_tc_tma_epilogue(int on) { g_tma_epilogue = on; }

static int epilogue_mode(const TcEpilogue& ep, int P, int N, bool train) {
    const bool lean_ok = !train && g_lean_epilogue && N % 32 == 0 && ep.clamp_max <= 0.f;
    if (lean_ok && ep.planes && !ep.c && !ep.residual && ep.alpha == 1.f && ep.n_planes == P && ep.ldp == N) return 1;
    if (lean_ok && ep.c && !ep.planes && ep.act == 0 && (ep.ldc & 3) == 0 && (!ep.residual || (ep.ldr & 3) == 0) &&
        ((uintptr_t)ep.c & 15) == 0 && ((uintptr_t)ep.residual & 15) == 0) return 2;
    return 0;
}

// CTA-pair launch: 256 x 256 tiles over pairs of SMs (N % 256 == 0, inference epilogues)
template <int P, int BK>
static int launch_cfg2(const PlaneMat& A, const PlaneMat& W, long M, int N, int K, const TcEpilogue& ep, cudaStream_t st) {
    using C = Cfg2<P, BK>;
    CUtensorMap ta, tb;
    int rc;
    if ((rc = make_map(&ta, A.ptr, M, K, A.ld, A.plane_stride, P, BM, BK))) return rc;
    if ((rc = make_map(&tb, W.ptr, N, K, W.ld, W.plane_stride, P, C::BN / 2, BK))) return rc;
    int em = epilogue_mode(ep, P, N, false);
    CUtensorMap tc = ta;
    static const bool tma_env = !(getenv("CSS_B200_GEMM_TMA_EPI") && getenv("CSS_B200_GEMM_TMA_EPI")[0] == '0');
    if (em == 1 && g_tma_epilogue && tma_env && ((uintptr_t)ep.bias & 15) == 0 && ((uintptr_t)ep.planes & 15) == 0) {
        // planes-only output: 32 x 32 tiles of each plane leave through the TMA store (EM 3)
        if ((rc = make_map(&tc, ep.planes, M, ep.ldp, ep.ldp, ep.plane_stride, P, 32, 32))) return rc;
        em = 3;
    }
    if (ep.ln_g) {
        CSS_REQUIRE(N == C::BN && ep.c && ep.planes && ep.ln_b && ep.act == 0 && ep.clamp_max <= 0.f && ep.n_planes == P && ep.ldp == N &&
                    (ep.ldc & 3) == 0 && (!ep.residual || (ep.ldr & 3) == 0) && ((uintptr_t)ep.c & 15) == 0 && ((uintptr_t)ep.residual & 15) == 0,
                    "gemm_tc: fused LayerNorm needs N == 256, fp32 + %d-plane outputs, no activation", P);
        em = 4;
    }
    static CssPerDevice attr;
    if (attr.first()) {
        cudaError_t e = cudaFuncSetAttribute(gemm_tc2_kernel<P, BK, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc2_kernel<P, BK, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc2_kernel<P, BK, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc2_kernel<P, BK, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc2_kernel<P, BK, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        CSS_REQUIRE(e == cudaSuccess, "gemm_tc: cannot reserve %d B of shared memory: %s", C::SMEM, cudaGetErrorString(e));
    }
    const long tiles = ((M + 2 * BM - 1) / (2 * BM)) * (N / C::BN);
    const int pairs = (int)(tiles < sm_count() / 2 ? tiles : sm_count() / 2);
    if (em == 4) gemm_tc2_kernel<P, BK, 4><<<2 * pairs, 384, C::SMEM, st>>>(ta, tb, tc, M, N, K, ep);
    else if (em == 3) gemm_tc2_kernel<P, BK, 3><<<2 * pairs, 384, C::SMEM, st>>>(ta, tb, tc, M, N, K, ep);
    else if (em == 1) gemm_tc2_kernel<P, BK, 1><<<2 * pairs, 384, C::SMEM, st>>>(ta, tb, tc, M, N, K, ep);
    else if (em == 2) gemm_tc2_kernel<P, BK, 2><<<2 * pairs, 384, C::SMEM, st>>>(ta, tb, tc, M, N, K, ep);
    else gemm_tc2_kernel<P, BK, 0><<<2 * pairs, 384, C::SMEM, st>>>(ta, tb, tc, M, N, K, ep);
    return css_check_launch("gemm_tc (CTA pair)");
}

template <int P, int BN, int BK>
static int launch_cfg(const PlaneMat& A, const PlaneMat& W, long M, int N, int K, const TcEpilogue& ep, cudaStream_t st, int mn = 0,
                      long k_valid = 0) {
    using C = Cfg<P, BN, BK>;
    CUtensorMap ta, tb;
    int rc;
    // K-major operand: (rows, K) matrix, box BK x tile rows.  MN-major operand: (K, rows) matrix, box 64 columns x BK rows.
    const long kv = k_valid > 0 ? k_valid : K;
    if (mn & 2) rc = make_map(&ta, A.ptr, kv, (int)M, A.ld, A.plane_stride, P, BK, 64);
    else rc = make_map(&ta, A.ptr, M, K, A.ld, A.plane_stride, P, BM, BK);
    if (rc) return rc;
    if (mn & 1) rc = make_map(&tb, W.ptr, kv, N, W.ld, W.plane_stride, P, BK, 64);
    else rc = make_map(&tb, W.ptr, N, K, W.ld, W.plane_stride, P, BN, BK);
    if (rc) return rc;
    const bool train = ep.mask != nullptr || ep.ksplit > 1 || mn != 0;
    const int em = epilogue_mode(ep, P, N, train);
    static CssPerDevice attr;
    if (attr.first()) {
        cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<P, BN, BK, false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc_kernel<P, BN, BK, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc_kernel<P, BN, BK, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc_kernel<P, BN, BK, true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if constexpr (P <= 2 && BN == 256) {
            if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc_kernel<P, BN, BK, true, 0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc_kernel<P, BN, BK, true, 0, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        }
        CSS_REQUIRE(e == cudaSuccess, "gemm_tc: cannot reserve %d B of shared memory: %s", C::SMEM, cudaGetErrorString(e));
    }
    const long tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN) * (ep.ksplit > 1 ? ep.ksplit : 1);
    const int grid = (int)(tiles < sm_count() ? tiles : sm_count());
    if (mn) {
        if constexpr (P <= 2 && BN == 256) {
            if (mn == 1) gemm_tc_kernel<P, BN, BK, true, 0, 1><<<grid, 384, C::SMEM, st>>>(ta, tb, M, N, K, ep);
            else gemm_tc_kernel<P, BN, BK, true, 0, 3><<<grid, 384, C::SMEM, st>>>(ta, tb, M, N, K, ep);
            return css_check_launch("gemm_tc (MN-major operands)");
        } else {
            CSS_REQUIRE(false, "gemm_tc: MN-major operands are built for 1 or 2 planes and 256-column tiles");
        }
    }
    if (train) gemm_tc_kernel<P, BN, BK, true, 0><<<grid, 384, C::SMEM, st>>>(ta, tb, M, N, K, ep);
    else if (em == 1) gemm_tc_kernel<P, BN, BK, false, 1><<<grid, 384, C::SMEM, st>>>(ta, tb, M, N, K, ep);
    else if (em == 2) gemm_tc_kernel<P, BN, BK, false, 2><<<grid, 384, C::SMEM, st>>>(ta, tb, M, N, K, ep);
    else gemm_tc_kernel<P, BN, BK, false, 0><<<grid, 384, C::SMEM, st>>>(ta, tb, M, N, K, ep);
    return css_check_launch("gemm_tc");
}

static int g_ln_fusion = 1;       // LayerNorm folded into the FFN-out GEMM epilogue (A/B switch; CSS_B200_GEMM_LN=0 disables)
extern "C" void css_gemm_tc_ln_fusion(int on) { g_ln_fusion = on; }
bool gemm_tc_ln_fusable(long M, int N, int P) {
    static const bool env = !(getenv("CSS_B200_GEMM_LN") && getenv("CSS_B200_GEMM_LN")[0] == '0');
    return g_ln_fusion && env && N == 256 && P <= 2 && M > 2 * BM && sm_count() >= 2 && two_cta_enabled();
}

static int g_tc_variant = 0;   // 0: BK=64 / 128B swizzle (default: 5-7 % faster over the 16-block network since the round-2 epilogues), 1: BK=32 / 64B swizzle (deeper ring)
extern "C" void css_gemm_tc_variant(int v) { g_tc_variant = v; }

int gemm_tc_planes(const PlaneMat& A, const PlaneMat& W, int P, long M, int N, int K, const GemmEpilogue& e, float* C, int ldc,
                   const PlaneMat* out, cudaStream_t st) {
    CSS_REQUIRE(P >= 1 && P <= 3, "gemm_tc: %d planes unsupported", P);
    CSS_REQUIRE(A.planes >= P && W.planes >= P, "gemm_tc: operands carry %d/%d planes, need %d", A.planes, W.planes, P);
    const int mn = (e.a_mn ? 2 : 0) | (e.b_mn ? 1 : 0);
    CSS_REQUIRE(mn != 2, "gemm_tc: an MN-major A needs an MN-major W as well (the dW form)");
    CSS_REQUIRE(M > 0 && N > 0 && A.ld >= (e.a_mn ? M : K) && W.ld >= (e.b_mn ? N : K), "gemm_tc: bad shape M=%ld N=%d K=%d", M, N, K);
    CSS_REQUIRE(!mn || (P <= 2 && N > 128 && !out && !e.ln_g), "gemm_tc: MN-major operands: 1-2 planes, N > 128, fp32 output");
    CSS_REQUIRE(C || out, "gemm_tc: no output");
    CSS_REQUIRE(K % KPAD == 0 && K > 0, "gemm_tc: K=%d must be a positive multiple of %d", K, KPAD);
    TcEpilogue ep = {e.bias, e.act, e.alpha, e.residual, e.ldr, e.clamp_max, C, ldc,
                     out ? reinterpret_cast<__nv_bfloat16*>(out->ptr) : nullptr, out ? out->plane_stride : 0,
                     out ? out->ld : 0, out ? out->planes : 0, reinterpret_cast<const __nv_bfloat16*>(e.relu_mask), e.ldmask, 1};
    if (e.accumulate) {
        CSS_REQUIRE(C && !out && !e.bias && !e.act && !e.residual && !e.relu_mask, "gemm_tc: accumulate mode takes a plain fp32 output");
        const int bn = N > 128 ? 256 : 128, bk = (N > 128 && g_tc_variant == 1) ? 32 : 64;
        const long tiles = ((M + BM - 1) / BM) * ((N + bn - 1) / bn);
        const int kblocks = K / bk;
        int want = (int)(sm_count() / tiles);
        if (want < 1) want = 1;
        if (want > kblocks) want = kblocks;
        const int kper = (kblocks + want - 1) / want;
        ep.ksplit = (kblocks + kper - 1) / kper;   // every split owns at least one k-block
        if (ep.ksplit < 2) ep.ksplit = 2;           // force the atomic epilogue (second split may not be empty: kblocks >= 2 always holds for K >= 128)
        CSS_REQUIRE(kblocks >= 2, "gemm_tc: accumulate mode needs K >= 128");
    }
    const bool wide = N > 128;
    const int v = g_tc_variant;
    const bool train = ep.mask != nullptr || ep.ksplit > 1;
    ep.ln_g = e.ln_g;
    ep.ln_b = e.ln_b;
    ep.ln_to_c = e.ln_g ? e.ln_to_c : 0;
    ep.ln2_g = (e.ln_g && e.ln_to_c) ? e.ln2_g : nullptr;
    ep.ln2_b = ep.ln2_g ? e.ln2_b : nullptr;
    CSS_REQUIRE(!ep.ln2_g || ep.ln2_b, "gemm_tc: ln2_g without ln2_b");
    CSS_REQUIRE(!ep.ln_to_c || (((uintptr_t)e.ln_g | (uintptr_t)e.ln_b) & 15) == 0, "gemm_tc: LayerNorm parameters must be 16-byte aligned");
    CSS_REQUIRE(!e.ln_g || (!train && gemm_tc_ln_fusable(M, N, P)), "gemm_tc: fused LayerNorm unsupported for M=%ld N=%d P=%d", M, N, P);
    if (mn) {
        if (P == 1) return v == 1 ? launch_cfg<1, 256, 32>(A, W, M, N, K, ep, st, mn, e.k_valid) : launch_cfg<1, 256, 64>(A, W, M, N, K, ep, st, mn, e.k_valid);
        return v == 1 ? launch_cfg<2, 256, 32>(A, W, M, N, K, ep, st, mn, e.k_valid) : launch_cfg<2, 256, 64>(A, W, M, N, K, ep, st, mn, e.k_valid);
    }
    if (!train && P <= 2 && N % 256 == 0 && M > 2 * BM && sm_count() >= 2 && two_cta_enabled()) {
        if (P == 1) return v == 1 ? launch_cfg2<1, 32>(A, W, M, N, K, ep, st) : launch_cfg2<1, 64>(A, W, M, N, K, ep, st);
        return v == 1 ? launch_cfg2<2, 32>(A, W, M, N, K, ep, st) : launch_cfg2<2, 64>(A, W, M, N, K, ep, st);
    }
    if (P == 1) {
        if (!wide) return launch_cfg<1, 128, 64>(A, W, M, N, K, ep, st);
        return v == 1 ? launch_cfg<1, 256, 32>(A, W, M, N, K, ep, st) : launch_cfg<1, 256, 64>(A, W, M, N, K, ep, st);
    }
    if (P == 2) {
        if (!wide) return launch_cfg<2, 128, 64>(A, W, M, N, K, ep, st);
        return v == 1 ? launch_cfg<2, 256, 32>(A, W, M, N, K, ep, st) : launch_cfg<2, 256, 64>(A, W, M, N, K, ep, st);
    }
    return v == 1 ? launch_cfg<3, 256, 32>(A, W, M, N, K, ep, st) : launch_cfg<3, 128, 64>(A, W, M, N, K, ep, st);
}

// fp32 (rows, cols; pitch ld) -> n_planes bf16 planes with pitch ldp; columns [cols, ldp) are zero filled
__global__ void __launch_bounds__(256)
split_planes_kernel(const float* __restrict__ x, long rows, int cols, int ld, __nv_bfloat16* __restrict__ out, long plane_stride,
                    int ldp, int n_planes) {
    const int pairs = ldp >> 1;
    const long total = rows * pairs;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long r = i / pairs;
        const int c = (int)(i - r * pairs) * 2;
        float a = c < cols ? x[r * ld + c] : 0.f;
        float b = c + 1 < cols ? x[r * ld + c + 1] : 0.f;
        for (int p = 0; p < n_planes; ++p) {
            const __nv_bfloat16 ha = __float2bfloat16_rn(a), hb = __float2bfloat16_rn(b);
            a -= __bfloat162float(ha);
            b -= __bfloat162float(hb);
            __nv_bfloat162 v;
            v.x = ha; v.y = hb;
            *reinterpret_cast<__nv_bfloat162*>(out + p * plane_stride + r * ldp + c) = v;
        }
    }
}

int split_planes_launch(const float* x, long rows, int cols, int ld, const PlaneMat& out, cudaStream_t st) {
    CSS_REQUIRE(rows > 0 && cols > 0 && out.ld >= cols && (out.ld & 1) == 0, "split_planes: bad shape");
    const long total = rows * (out.ld >> 1);
    long grid = (total + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    split_planes_kernel<<<(unsigned)grid, 256, 0, st>>>(x, rows, cols, ld, reinterpret_cast<__nv_bfloat16*>(out.ptr),
                                                       out.plane_stride, out.ld, out.planes);
    return css_check_launch("split_planes");
}

// Public test/utility entry: C = epilogue(A . W^T) on the tcgen05 path from fp32 operands.  The operands are split
// into `planes` bf16 planes in `scratch` first (the network path keeps its activations in plane form instead).
extern "C" size_t css_gemm_tc_scratch_bytes(long M, int N, int K, int planes) {
    const size_t kp = (size_t)((K + KPAD - 1) / KPAD) * KPAD;
    return (size_t)planes * ((size_t)M + N) * kp * 2 + 512;
}
extern "C" int css_gemm_tc(const float* A, int lda, const float* W, int ldw, const float* bias, float* C, int ldc, long M,
                           int N, int K, int act, float alpha, const float* residual, int ldr, float clamp_max, int planes,
                           void* scratch, size_t scratch_bytes, cudaStream_t st) {
    CSS_REQUIRE(planes >= 1 && planes <= 3, "css_gemm_tc: planes must be 1..3");
    CSS_REQUIRE(M > 0 && N > 0 && K > 0, "css_gemm_tc: empty shape");
    CSS_REQUIRE(scratch && scratch_bytes >= css_gemm_tc_scratch_bytes(M, N, K, planes), "css_gemm_tc: scratch too small");
    const int kp = (K + KPAD - 1) / KPAD * KPAD;
    char* base = reinterpret_cast<char*>(((uintptr_t)scratch + 255) & ~(uintptr_t)255);
    PlaneMat a = {base, M * (long)kp, kp, planes};
    PlaneMat w = {base + (size_t)planes * M * kp * 2, (long)N * kp, kp, planes};
    int rc;
    if ((rc = split_planes_launch(A, M, K, lda, a, st))) return rc;
    if ((rc = split_planes_launch(W, N, K, ldw, w, st))) return rc;
    GemmEpilogue e = {bias, act, alpha, residual, ldr, clamp_max};
    css_prof_begin(0, 2.0 * (double)M * N * K, st);
    rc = gemm_tc_planes(a, w, planes, M, N, kp, e, C, ldc, nullptr, st);
    css_prof_end(st);
    return rc;
}

// Test entry for the MN-major operand forms (training GEMMs): mode 1: C = A . W^T with W handed over as Wt = W^T (K, N);
// mode 3: additionally A handed over as At = A^T (K, M), result ACCUMULATED into C (split-K + atomics, the dW form).
// scratch >= 2 * K * (pitch of M + pitch of N) * 2 bytes + 512.
extern "C" int css_gemm_tc_mn(const float* A, int lda, const float* Wt, int ldw, float* C, int ldc, long M, int N, int K, int mode,
                              int planes, void* scratch, size_t scratch_bytes, cudaStream_t st) {
    CSS_REQUIRE((mode == 1 || mode == 3) && planes >= 1 && planes <= 2 && K % KPAD == 0, "css_gemm_tc_mn: bad mode / planes / K");
    const int mp = (int)((M + 63) / 64 * 64), np = (N + 63) / 64 * 64;
    char* base = reinterpret_cast<char*>(((uintptr_t)scratch + 255) & ~(uintptr_t)255);
    PlaneMat a, w;
    size_t need;
    int rc;
    if (mode == 3) {
        a = {base, (long)K * mp, mp, planes};
        need = (size_t)planes * K * mp * 2;
        if ((rc = split_planes_launch(A, K, (int)M, lda, a, st))) return rc;
    } else {
        a = {base, M * (long)K, K, planes};
        need = (size_t)planes * M * K * 2;
        if ((rc = split_planes_launch(A, M, K, lda, a, st))) return rc;
    }
    need = (need + 255) & ~(size_t)255;
    w = {base + need, (long)K * np, np, planes};
    CSS_REQUIRE(scratch_bytes >= need + (size_t)planes * K * np * 2 + 512, "css_gemm_tc_mn: scratch too small");
    if ((rc = split_planes_launch(Wt, K, N, ldw, w, st))) return rc;
    GemmEpilogue e = {nullptr, 0, 1.f, nullptr, 0, 0.f};
    e.b_mn = 1;
    e.a_mn = mode == 3;
    e.accumulate = mode == 3;
    return gemm_tc_planes(a, w, planes, M, N, K, e, C, ldc, nullptr, st);
}
