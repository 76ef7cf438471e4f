// prims.cu -- one kernel per cdsp static (include/cdsp.h:36-71 in the reference tree), for
// code that calls cdsp:: directly on host buffers.  Not the fast path: the fused kernels in
// fft_kernels.cu / apply_kernel.cu never call these.  Each kernel states the volk / fftw call
// it replaces.
#include "coherent_cuda.h"
#include "kernels.h"
#include <string>
#include <vector>

namespace coh {

constexpr int kT = 256;
static inline int grid_for(int n) { int g = (n + kT - 1) / kT; return g < 1 ? 1 : (g > 1184 ? 1184 : g); }

// cdsp::convtosigned (cdsp.cc:21-34): 8 bytes at a time, n>>3 iterations
__global__ void k_convtosigned(const uint8_t *in, uint8_t *out, int n8) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += gridDim.x * blockDim.x)
        reinterpret_cast<unsigned long long *>(out)[i] =
            reinterpret_cast<const unsigned long long *>(in)[i] ^ 0x8080808080808080ull;
}
// volk_8i_s32f_convert_32f (cdsp.cc:36-44)
__global__ void k_convtofloat(float *out, const int8_t *in, int n) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        out[i] = __fmul_rn((float)in[i], 0x1.020408p-7f);
}
// volk_32f_s32f_convert_8i (cdsp.cc:51-54)
__global__ void k_convto8bit(int8_t *out, const float *in, int n) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float v = __fmul_rn(in[i], 127.0f);
        v = (v != v) ? 0.0f : fminf(fmaxf(v, -128.0f), 127.0f);
        out[i] = (int8_t)__float2int_rn(v);
    }
}
// volk_32fc_s32fc_multiply_32fc (cdsp.cc:46-49), AVX+FMA kernel rounding
__global__ void k_scalarmul(float2 *out, const float2 *in, float2 s, int n) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float2 a = in[i];
        out[i] = make_float2(__fmaf_rn(a.x, s.x, -__fmul_rn(a.y, s.y)), __fmaf_rn(a.y, s.x, __fmul_rn(a.x, s.y)));
    }
}
// volk_32fc_magnitude_squared_32f (cdsp.cc:100-103)
__global__ void k_magsquared(float *out, const float2 *in, int n) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float2 a = in[i];
        out[i] = __fadd_rn(__fmul_rn(a.x, a.x), __fmul_rn(a.y, a.y));
    }
}
// volk_32fc_x2_multiply_conjugate_32fc (cdsp.cc:105-108)
__global__ void k_conjugatemul(float2 *out, const float2 *a, const float2 *b, int n) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float2 x = a[i], y = b[i];
        out[i] = make_float2(__fadd_rn(__fmul_rn(x.x, y.x), __fmul_rn(x.y, y.y)),
                             __fsub_rn(__fmul_rn(x.y, y.x), __fmul_rn(x.x, y.y)));
    }
}
// volk_32fc_x2_conjugate_dot_prod_32fc (cdsp.cc:61-66): single CTA, double accumulation
__global__ void k_conj_dot(float2 *res, const float2 *a, const float2 *b, int n) {
    __shared__ double sr[kT / 32], si[kT / 32];
    double re = 0, im = 0;
    for (int i = threadIdx.x; i < n; i += kT) {
        float2 x = a[i], y = b[i];
        re += (double)x.x * y.x + (double)x.y * y.y;
        im += (double)x.y * y.x - (double)x.x * y.y;
    }
    for (int o = 16; o; o >>= 1) { re += __shfl_down_sync(~0u, re, o); im += __shfl_down_sync(~0u, im, o); }
    if ((threadIdx.x & 31) == 0) { sr[threadIdx.x >> 5] = re; si[threadIdx.x >> 5] = im; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kT / 32; w++) { re += sr[w]; im += si[w]; }
        *res = make_float2((float)re, (float)im);
    }
}
// volk_32f_x2_dot_prod_32f(in,in) (cdsp.cc:68-73): sum of squares
__global__ void k_sumsq(float *res, const float *in, int n) {
    __shared__ double s[kT / 32];
    double acc = 0;
    for (int i = threadIdx.x; i < n; i += kT) acc += (double)in[i] * in[i];
    for (int o = 16; o; o >>= 1) acc += __shfl_down_sync(~0u, acc, o);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) { for (int w = 1; w < kT / 32; w++) acc += s[w]; *res = (float)acc; }
}
// volk_32f_index_max_32u (cdsp.cc:135-139): first index of the maximum
__global__ void k_indexofmax(uint32_t *idx, const float *in, int n) {
    __shared__ float sv[kT / 32]; __shared__ uint32_t si[kT / 32];
    float best = -INFINITY; uint32_t bi = 0xffffffffu;
    for (int i = threadIdx.x; i < n; i += kT) { float v = in[i]; if (v > best) { best = v; bi = i; } }
    for (int o = 16; o; o >>= 1) {
        float ov = __shfl_down_sync(~0u, best, o); uint32_t oi = __shfl_down_sync(~0u, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kT / 32; w++) if (sv[w] > best || (sv[w] == best && si[w] < bi)) { best = sv[w]; bi = si[w]; }
        *idx = bi == 0xffffffffu ? 0 : bi;
    }
}

// fftwf_execute_dft on a howmany-batched plan (ccoherent.cc:87-93; cdsp.cc:110-120): in-place
// radix-2 decimation-in-frequency in shared memory, bit-reversed write-back.  One CTA per
// transform, n <= 16384.
__global__ void __launch_bounds__(512, 1) k_fft_generic(float2 *out, const float2 *in, int log2n, int sign) {
    extern __shared__ __align__(16) float2 sm[];
    const int n = 1 << log2n;
    const float2 *x = in + (size_t)blockIdx.x * n;
    float2 *y = out + (size_t)blockIdx.x * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) sm[i] = x[i];
    __syncthreads();
    for (int s = log2n - 1; s >= 0; s--) {
        const int half = 1 << s;
        for (int t = threadIdx.x; t < n / 2; t += blockDim.x) {
            int lo = t & (half - 1), i0 = ((t >> s) << (s + 1)) + lo, i1 = i0 + half;
            float2 a = sm[i0], b = sm[i1];
            float sn, cs;
            sincospif((float)sign * (float)lo / (float)half, &sn, &cs);
            float2 d = make_float2(a.x - b.x, a.y - b.y);
            sm[i0] = make_float2(a.x + b.x, a.y + b.y);
            sm[i1] = make_float2(d.x * cs - d.y * sn, d.x * sn + d.y * cs);
        }
        __syncthreads();
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x) y[__brev((unsigned)i) >> (32 - log2n)] = sm[i];
}

#define LAUNCH1(k, n, ...) k<<<grid_for(n), kT, 0, st>>>(__VA_ARGS__); return cudaGetLastError()
cudaError_t prim_convtosigned(const uint8_t *in, uint8_t *out, int n, cudaStream_t st) { LAUNCH1(k_convtosigned, n >> 3, in, out, n >> 3); }
cudaError_t prim_convtofloat(float *out, const int8_t *in, int n, cudaStream_t st) { LAUNCH1(k_convtofloat, n, out, in, n); }
cudaError_t prim_convto8bit(int8_t *out, const float *in, int n, cudaStream_t st) { LAUNCH1(k_convto8bit, n, out, in, n); }
cudaError_t prim_scalarmul(float2 *out, const float2 *in, float2 s, int n, cudaStream_t st) { LAUNCH1(k_scalarmul, n, out, in, s, n); }
cudaError_t prim_magsquared(float *out, const float2 *in, int n, cudaStream_t st) { LAUNCH1(k_magsquared, n, out, in, n); }
cudaError_t prim_conjugatemul(float2 *out, const float2 *a, const float2 *b, int n, cudaStream_t st) { LAUNCH1(k_conjugatemul, n, out, a, b, n); }
cudaError_t prim_conj_dot(float2 *res, const float2 *a, const float2 *b, int n, cudaStream_t st) { k_conj_dot<<<1, kT, 0, st>>>(res, a, b, n); return cudaGetLastError(); }
cudaError_t prim_sumsq(float *res, const float *in, int n, cudaStream_t st) { k_sumsq<<<1, kT, 0, st>>>(res, in, n); return cudaGetLastError(); }
cudaError_t prim_indexofmax(uint32_t *idx, const float *in, int n, cudaStream_t st) { k_indexofmax<<<1, kT, 0, st>>>(idx, in, n); return cudaGetLastError(); }
bool prim_fft_supported(int n) { return n >= 2 && n <= 16384 && (n & (n - 1)) == 0; }
cudaError_t prim_fft(float2 *out, const float2 *in, int n, int howmany, int sign, const float2 *, cudaStream_t st) {
    int log2n = 0; while ((1 << log2n) < n) log2n++;
    cudaError_t e = cudaFuncSetAttribute(k_fft_generic, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 8);
    if (e != cudaSuccess) return e;
    k_fft_generic<<<howmany, 512, (size_t)n * 8, st>>>(out, in, log2n, sign);
    return cudaGetLastError();
}

}  // namespace coh
