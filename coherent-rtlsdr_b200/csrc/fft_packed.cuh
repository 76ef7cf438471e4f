// fft_packed.cuh -- two complex transforms at once on Blackwell's packed fp32 pipe.
//
// The correlation kernels transform two L-point sequences with identical structure and
// twiddles (the even-bin half "A" and the odd-bin half "B" of the zero-padded 2L-point
// transform, see fft_kernels.cu).  A PC holds one element of each: r = (A.re, B.re),
// i = (A.im, B.im), so that every butterfly add is ONE FADD2 and every twiddle product four
// FMUL2/FFMA2 for both halves (sm_100a packed f32x2; the second source of FFMA2/FMUL2 may be a
// scalar register or immediate broadcast to both lanes, so scalar twiddles cost no extra moves).
// In shared memory an element is one float4 {A.re, B.re, A.im, B.im}: all traffic is 16-byte
// LDS/STS, which on this part moves 126 B/clk/SM against 63 B/clk/SM for 8-byte accesses
// (tools/ubench_fma.cu).
//
// Stands in for fftwf_execute_dft as called at src/ccoherent.cc:174,182 of the reference
// (sign -1 forward, +1 backward, unnormalised); textbook decimation on compile-time constants.
#pragma once
#include <cuda_runtime.h>
#include "w32_consts.cuh"

namespace coh {

struct PC { float2 r, i; };

__device__ __forceinline__ float2 splat(float v) { return make_float2(v, v); }
__device__ __forceinline__ float2 neg2(float2 v) { return make_float2(-v.x, -v.y); }

__device__ __forceinline__ PC pc_from(float4 v) { PC p; p.r = make_float2(v.x, v.y); p.i = make_float2(v.z, v.w); return p; }
__device__ __forceinline__ float4 pc_to(PC p) { return make_float4(p.r.x, p.r.y, p.i.x, p.i.y); }

__device__ __forceinline__ PC pc_add(PC a, PC b) { PC o; o.r = __fadd2_rn(a.r, b.r); o.i = __fadd2_rn(a.i, b.i); return o; }
__device__ __forceinline__ PC pc_sub(PC a, PC b) { PC o; o.r = __fadd2_rn(a.r, neg2(b.r)); o.i = __fadd2_rn(a.i, neg2(b.i)); return o; }

// both lanes times the same scalar complex w
__device__ __forceinline__ PC pc_mul(PC a, float2 w) {
    PC o;
    // o.r accumulates onto a.r's product and o.i onto a.i's so that each result can stay in its
    // operand's register pair: the {r, i} register quad an LDS.128 delivered is still a quad
    // when STS.128 needs it (otherwise ptxas spends four MOVs per store)
    const float2 t1 = __fmul2_rn(a.i, splat(-w.y)), t2 = __fmul2_rn(a.r, splat(w.y));
    o.r = __ffma2_rn(a.r, splat(w.x), t1);
    o.i = __ffma2_rn(a.i, splat(w.x), t2);
    return o;
}
// lane-wise product with a packed pair of twiddles t = (tA, tB)
__device__ __forceinline__ PC pc_mulp(PC a, PC t) {
    PC o;
    const float2 t1 = __fmul2_rn(a.i, neg2(t.i)), t2 = __fmul2_rn(a.r, t.i);
    o.r = __ffma2_rn(a.r, t.r, t1);
    o.i = __ffma2_rn(a.i, t.r, t2);
    return o;
}
// lane-wise a * conj(t)   (cdsp::conjugatemul, src/cdsp.cc:105-108)
__device__ __forceinline__ PC pc_mulcp(PC a, PC t) {
    PC o;
    const float2 t1 = __fmul2_rn(a.i, t.i), t2 = __fmul2_rn(neg2(a.r), t.i);
    o.r = __ffma2_rn(a.r, t.r, t1);
    o.i = __ffma2_rn(a.i, t.r, t2);
    return o;
}

// a * exp(DIR * 2*pi*i * M/32), M a compile-time constant after unrolling
template <int DIR>
__device__ __forceinline__ PC pc_mul_w32(PC a, int M) {
    M &= 31;
    PC o;
    if (M == 0)  return a;
    if (M == 16) { o.r = neg2(a.r); o.i = neg2(a.i); return o; }
    if (M == 8)  { if (DIR < 0) { o.r = a.i; o.i = neg2(a.r); } else { o.r = neg2(a.i); o.i = a.r; } return o; }
    if (M == 24) { if (DIR < 0) { o.r = neg2(a.i); o.i = a.r; } else { o.r = a.i; o.i = neg2(a.r); } return o; }
    const float c = kW32c[M];
    const float s = DIR < 0 ? -kW32s[M] : kW32s[M];
    return pc_mul(a, make_float2(c, s));
}

// In-place, natural-order DFT of R packed points held in registers.
template <int R, int DIR> struct PDft {
    static __device__ __forceinline__ void run(PC (&v)[R]) {
        PC e[R / 2], o[R / 2];
#pragma unroll
        for (int k = 0; k < R / 2; k++) { e[k] = v[2 * k]; o[k] = v[2 * k + 1]; }
        PDft<R / 2, DIR>::run(e);
        PDft<R / 2, DIR>::run(o);
#pragma unroll
        for (int k = 0; k < R / 2; k++) {
            PC t = pc_mul_w32<DIR>(o[k], k * (32 / R));
            v[k] = pc_add(e[k], t);
            v[k + R / 2] = pc_sub(e[k], t);
        }
    }
};
template <int DIR> struct PDft<1, DIR> {
    static __device__ __forceinline__ void run(PC (&)[1]) {}
};

// packed twiddle pairs T[k] = t0 * w^k, k = 0..R-1 (both lanes advance by the same scalar w);
// multiplication depth <= log2(R)
template <int R>
__device__ __forceinline__ void twiddle_pairs(PC t0, float2 w, PC (&T)[R]) {
    float2 wp[5];                                   // w^1, w^2, w^4, w^8, w^16
    wp[0] = w;
#pragma unroll
    for (int j = 1; j < 5; j++)
        wp[j] = make_float2(wp[j - 1].x * wp[j - 1].x - wp[j - 1].y * wp[j - 1].y, 2.0f * wp[j - 1].x * wp[j - 1].y);
    T[0] = t0;
#pragma unroll
    for (int k = 1; k < R; k++) {
        int p = 1, j = 0;
        while (p * 2 <= k) { p *= 2; j++; }
        T[k] = pc_mul(T[k - p], wp[j]);
    }
}

// shared-memory padding, in float4 elements: one pad element after every 16 keeps the
// stride-2^j column accesses of all passes on distinct 16-byte bank groups
__host__ __device__ __forceinline__ int ppad(int a) { return a + (a >> 4); }
__host__ __device__ constexpr int ppadded_len(int n) { return n + (n >> 4) + 1; }
// the same with the padding distance as a parameter: PADSH = 4 is ppad(); PADSH = 5 (one pad element per 32)
// is what the fused 32-point stage of the persistent L = 8192 kernel needs (fft_passes.cuh, fused_b_stage)
template <int PADSH> __host__ __device__ __forceinline__ int ppad_t(int a) { return a + (a >> PADSH); }
template <int PADSH> __host__ __device__ constexpr int ppadded_len_t(int n) { return n + (n >> PADSH) + 1; }

}  // namespace coh
