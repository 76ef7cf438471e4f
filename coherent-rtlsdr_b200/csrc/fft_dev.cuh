// fft_dev.cuh -- register-resident radix-2/4/8/16 DFT butterflies and twiddle helpers used by
// the correlation kernels (fft_kernels.cu) and the generic cdsp::fft kernel.
//
// Stands in for fftwf_execute_dft as called at src/ccoherent.cc:174,182 (reference tree):
// sign -1 forward, +1 backward, both unnormalised.  Nothing here is derived from FFTW code;
// it is a textbook decimation split evaluated on compile-time constants.
#pragma once
#include <cuda_runtime.h>
#include "w32_consts.cuh"

namespace coh {

__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cmulc(float2 a, float2 b) {           // a * conj(b)
    return make_float2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y);
}
__device__ __forceinline__ float2 cconj(float2 a) { return make_float2(a.x, -a.y); }

// a * exp(DIR * 2*pi*i * M/32) with M a compile-time constant after unrolling
template <int DIR>
__device__ __forceinline__ float2 mul_w32(float2 a, int M) {
    M &= 31;
    if (M == 0)  return a;
    if (M == 16) return make_float2(-a.x, -a.y);
    if (M == 8)  return DIR < 0 ? make_float2(a.y, -a.x) : make_float2(-a.y, a.x);
    if (M == 24) return DIR < 0 ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x);
    const float c = kW32c[M];
    const float s = DIR < 0 ? -kW32s[M] : kW32s[M];
    return make_float2(a.x * c - a.y * s, a.x * s + a.y * c);
}

// In-place, natural-order DFT of R points held in registers.
template <int R, int DIR> struct Dft {
    static __device__ __forceinline__ void run(float2 (&v)[R]) {
        float2 e[R / 2], o[R / 2];
#pragma unroll
        for (int i = 0; i < R / 2; i++) { e[i] = v[2 * i]; o[i] = v[2 * i + 1]; }
        Dft<R / 2, DIR>::run(e);
        Dft<R / 2, DIR>::run(o);
#pragma unroll
        for (int k = 0; k < R / 2; k++) {
            float2 t = mul_w32<DIR>(o[k], k * (32 / R));
            v[k] = cadd(e[k], t);
            v[k + R / 2] = csub(e[k], t);
        }
    }
};
template <int DIR> struct Dft<1, DIR> {
    static __device__ __forceinline__ void run(float2 (&)[1]) {}
};

// pw[k] = w^k for k = 0..R-1 with multiplication depth <= log2(R)
template <int R>
__device__ __forceinline__ void twiddle_powers(float2 w, float2 (&pw)[R]) {
    pw[0] = make_float2(1.0f, 0.0f);
    if (R > 1) pw[1] = w;
#pragma unroll
    for (int k = 2; k < R; k++) {
        // largest power of two <= k
        int p = 1;
        while (p * 2 <= k) p *= 2;
        pw[k] = (k == p) ? cmul(pw[p / 2], pw[p / 2]) : cmul(pw[p], pw[k - p]);
    }
}

// shared-memory padding: 2 complex words after every 32 keep stride-2^j column accesses of
// the in-place passes on distinct banks (DESIGN.md "shared-memory layout")
__host__ __device__ __forceinline__ int padidx(int a) { return a + ((a >> 5) << 1); }
__host__ __device__ constexpr int padded_len(int n) { return n + (n >> 4); }

}  // namespace coh
