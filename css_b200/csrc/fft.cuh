// Warp-level 512-point real FFT / inverse real FFT staged in shared memory.
//
// A 512-point real transform is done as one 256-point complex Stockham radix-4 FFT
// (4 passes, ping-pong between two per-warp shared buffers) plus the even/odd split.
// One warp owns one frame; every lane does two radix-4 butterflies per pass.
#pragma once
#include "common.cuh"

struct FftTables {
    float2 tw[512];   // tw[k] = exp(-2*pi*i*k/512)
    float win[512];   // periodic Hann window (torch.hann_window(512))
};

__device__ __forceinline__ void fft_tables_init(FftTables* t) {
    for (int k = threadIdx.x; k < 512; k += blockDim.x) {
        float s, c;
        sincospif((float)k / 256.0f, &s, &c);
        t->tw[k] = make_float2(c, -s);
        t->win[k] = 0.5f - 0.5f * c;
    }
}

// The twiddles of a lane's butterflies do not depend on the frame: passes 1-3 x two butterflies x three factors, kept in
// registers by the persistent warps (read from the shared table they were 4- to 8-way bank conflicts -- 4, 16 or 32 distinct
// addresses on a 256 / 64 / 16-byte stride -- and about a third of the kernel's shared-memory wavefronts).
struct FftLaneTw { float2 w[3][2][3]; };
__device__ __forceinline__ void fft_lane_twiddles(FftLaneTw& t, const float2* __restrict__ tw, int lane) {
#pragma unroll
    for (int pass = 1; pass < 4; ++pass) {
        const int Ns = 1 << (2 * pass), twstep = 128 >> (2 * pass);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int k = (lane + 32 * h) & (Ns - 1);
#pragma unroll
            for (int q = 0; q < 3; ++q) t.w[pass - 1][h][q] = tw[(q + 1) * k * twstep];
        }
    }
}

// Forward 256-point complex FFT of `a` (natural order in, natural order out).
// Result pointer is returned (it is `a` again after 4 passes).  Warp-synchronous.
__device__ __forceinline__ float2* warp_fft256(float2* a, float2* b, const FftLaneTw& ltw, int lane) {
    float2* in = a;
    float2* out = b;
#pragma unroll
    for (int pass = 0; pass < 4; ++pass) {
        const int Ns = 1 << (2 * pass);          // 1,4,16,64
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int j = lane + 32 * h;         // butterfly index 0..63
            const int k = j & (Ns - 1);
            float2 v0 = in[j], v1 = in[j + 64], v2 = in[j + 128], v3 = in[j + 192];
            if (pass > 0) {                       // pass 0: k = 0, all three factors are 1
                v1 = cmul(v1, ltw.w[pass > 0 ? pass - 1 : 0][h][0]);
                v2 = cmul(v2, ltw.w[pass > 0 ? pass - 1 : 0][h][1]);
                v3 = cmul(v3, ltw.w[pass > 0 ? pass - 1 : 0][h][2]);
            }
            float2 s02 = make_float2(v0.x + v2.x, v0.y + v2.y);
            float2 d02 = make_float2(v0.x - v2.x, v0.y - v2.y);
            float2 s13 = make_float2(v1.x + v3.x, v1.y + v3.y);
            float2 d13 = make_float2(v1.y - v3.y, v3.x - v1.x);   // -i * (v1 - v3)
            const int j0 = ((j - k) << 2) + k;
            if (pass == 0) {
                // Ns = 1: the four outputs of a butterfly are neighbours -> two 16-byte stores, 32 contiguous bytes per lane
                // (four 8-byte stores with a 32-byte stride between lanes were an 8-way bank conflict each; the kernel sits on
                // the shared-memory pipe)
                // (lanes 4-7, 12-15, ... store their halves in the opposite order, so that one store instruction covers all 32
                // banks with eight lanes: 4 wavefronts per instruction, the minimum for 512 bytes)
                float4* o4 = reinterpret_cast<float4*>(out + 4 * j);
                const float4 lo = make_float4(s02.x + s13.x, s02.y + s13.y, d02.x + d13.x, d02.y + d13.y);
                const float4 hi = make_float4(s02.x - s13.x, s02.y - s13.y, d02.x - d13.x, d02.y - d13.y);
                const int flip = (lane >> 2) & 1;
                o4[flip] = flip ? hi : lo;
                o4[flip ^ 1] = flip ? lo : hi;
                continue;
            }
            float2 o0 = make_float2(s02.x + s13.x, s02.y + s13.y), o1 = make_float2(d02.x + d13.x, d02.y + d13.y);
            float2 o2 = make_float2(s02.x - s13.x, s02.y - s13.y), o3 = make_float2(d02.x - d13.x, d02.y - d13.y);
            if (pass == 1) {
                // Ns = 4: a group of four lanes writes 32 contiguous bytes and the eight groups of a warp sit 128 bytes apart
                // -- the same banks for every output index.  Group g therefore stores output (s + g) mod 4 in store s (the
                // four outputs are rotated by g in registers): four groups cover the 32 banks, 2 wavefronts instead of 8.
                const int g = (lane >> 2) & 3;
                if (g & 1) { const float2 t = o0; o0 = o1; o1 = o2; o2 = o3; o3 = t; }
                if (g & 2) { float2 t = o0; o0 = o2; o2 = t; t = o1; o1 = o3; o3 = t; }
                out[j0 + ((g) & 3) * Ns] = o0;
                out[j0 + ((g + 1) & 3) * Ns] = o1;
                out[j0 + ((g + 2) & 3) * Ns] = o2;
                out[j0 + ((g + 3) & 3) * Ns] = o3;
                continue;
            }
            out[j0] = o0;
            out[j0 + Ns] = o1;
            out[j0 + 2 * Ns] = o2;
            out[j0 + 3 * Ns] = o3;
        }
        __syncwarp();
        float2* tmp = in; in = out; out = tmp;
    }
    return in;
}

// After warp_fft256 of z[n] = x[2n] + i x[2n+1]:  X[k], k = 0..256.
__device__ __forceinline__ float2 rfft_bin(const float2* Z, const float2* __restrict__ tw, int k) {
    float2 zk = Z[k & 255];
    float2 zc = Z[(256 - k) & 255];
    zc.y = -zc.y;                                              // conj(Z[256-k])
    float2 e = make_float2(0.5f * (zk.x + zc.x), 0.5f * (zk.y + zc.y));
    float2 o = make_float2(0.5f * (zk.y - zc.y), -0.5f * (zk.x - zc.x));   // -i/2 (zk - zc)
    float2 wo = cmul(o, tw[k]);
    return make_float2(e.x + wo.x, e.y + wo.y);
}

// Inverse: pack one-sided spectrum X[0..256] into conj(Z) so that a forward FFT followed by a
// conjugate gives z[n] = x[2n] + i x[2n+1] (times 256).  Imag of DC/Nyquist is ignored (c2r).
__device__ __forceinline__ float2 irfft_pack(float2 xk, float2 xc /* X[256-k] */, const float2* __restrict__ tw, int k) {
    // E = (X[k] + conj(X[256-k]))/2 ; O = (X[k] - conj(X[256-k]))/2 * conj(tw[k]) ; Z = E + iO
    float2 e = make_float2(0.5f * (xk.x + xc.x), 0.5f * (xk.y - xc.y));
    float2 d = make_float2(0.5f * (xk.x - xc.x), 0.5f * (xk.y + xc.y));
    float2 w = tw[k];
    float2 o = make_float2(d.x * w.x + d.y * w.y, d.y * w.x - d.x * w.y);   // d * conj(w)
    float2 z = make_float2(e.x - o.y, e.y + o.x);
    z.y = -z.y;                                                // conj for the forward-FFT trick
    return z;
}
