// fft_kernels.cu -- correlation kernels of the block sync path (sm_100a).
//
//   k_ref_spectrum : convert the reference-noise block and transform it           (K1)
//   k_correlate    : per (block, signal channel): convert, forward transform, multiply by the
//                    conjugate reference spectrum, inverse transform, |.|^2 arg-max  (K1+K2 fused)
//
// Replaces, in the reference tree: crtlsdr::convtofloat / crefsdr::convtofloat
// (src/crtlsdr.cc:205-218), ccoherent::queuelag's 128 KiB memcpy (src/ccoherent.cc:123-142) and
// ccoherent::computelag's fft -> conjugatemul -> fft -> magsquared -> indexofmax chain
// (src/ccoherent.cc:174-192).
//
// Transform layout.  The reference zero-pads to N = 2L (signal in [0,L), ref in [L,2L)).  Half
// of every input being zero, the N-point forward transform splits exactly into two L-point
// ones: even bins = FFT_L(x), odd bins = FFT_L(x * W_N^n); for the reference channel the
// half-block shift contributes (-1)^k.  The inverse needs all N outputs:
//   c[m] = A[m] + conj(W_N^m) B[m],  c[m+L] = A[m] - conj(W_N^m) B[m],  0 <= m < L,
// with A = IFFT_L(even bins), B = IFFT_L(odd bins).  Both halves ("A", "B") live side by side
// in shared memory and every pass processes both with the same twiddles.
// Forward passes are decimation-in-frequency (natural in, digit-reversed out), inverse passes
// decimation-in-time (digit-reversed in, natural out), so no reordering pass exists and the
// reference spectrum is stored in the same digit-reversed order.
#include "fft_dev.cuh"
#include "kernels.h"

namespace coh {

constexpr int kFftThreads = 512;

template <int LOG2L> struct Plan {
    static constexpr int L = 1 << LOG2L;
    static constexpr int S1 = L / 16;                       // stride of pass 1 (radix 16)
    static constexpr int M1 = L / 16;                       // sub-transform length after pass 1
    static constexpr int RA = (M1 / 2 >= 16) ? 16 : (M1 / 2);   // mid pass A radix (1 = none)
    static constexpr int SA = (RA > 1) ? M1 / RA : M1;
    static constexpr int RB = (RA > 1 && SA > 2) ? SA / 2 : 1;  // mid pass B radix, stride 2
    static constexpr int LP = padded_len(L);
    static constexpr size_t kSmemBytes = 2 * (size_t)LP * sizeof(float2);
};

__device__ __forceinline__ float2 load_sample(const uint8_t *x, int n, uint32_t xormask) {
    // cdsp::convtosigned (src/cdsp.cc:21-34) + volk_8i_s32f_convert_32f (src/cdsp.cc:41-44)
    unsigned short raw = reinterpret_cast<const unsigned short *>(x)[n];
    raw ^= (unsigned short)xormask;
    const float c = 0x1.020408p-7f;                 // float(1.0/127.0)
    float re = (float)(signed char)(raw & 0xff) * c;
    float im = (float)(signed char)(raw >> 8) * c;
    return make_float2(re, im);
}

// ---- pass 1: global bytes -> shared, radix 16 at stride L/16, both halves ------------------
template <int LOG2L, bool NEG_B>
__device__ __forceinline__ void pass1_from_global(const uint8_t *x, uint32_t xormask,
                                                  float2 *sA, float2 *sB, const float2 *__restrict__ TN) {
    using P = Plan<LOG2L>;
    constexpr int S = P::S1;
    for (int low = threadIdx.x; low < S; low += kFftThreads) {
        float2 a[16], b[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
            a[k] = load_sample(x, low + k * S, xormask);
            b[k] = mul_w32<-1>(a[k], k);              // W_N^(k*L/16) = W_32^k
        }
        Dft<16, -1>::run(a);
        Dft<16, -1>::run(b);
        float2 w = __ldg(&TN[2 * low]);               // W_L^low
        float2 g = __ldg(&TN[low]);                   // W_N^low (odd-bin modulation)
        if (NEG_B) g = make_float2(-g.x, -g.y);       // ref sits in [L,2L): odd bins change sign
        float2 pw[16];
        twiddle_powers<16>(w, pw);
#pragma unroll
        for (int k = 0; k < 16; k++) {
            float2 ta = (k == 0) ? a[0] : cmul(a[k], pw[k]);
            float2 tb = cmul(b[k], (k == 0) ? g : cmul(g, pw[k]));
            sA[padidx(low + k * S)] = ta;
            sB[padidx(low + k * S)] = tb;
        }
    }
}

// ---- in-place mid passes over both halves ---------------------------------------------------
template <int LOG2L, int R, int S, int DIR>
__device__ __forceinline__ void mid_pass(float2 *sA, float2 *sB, const float2 *__restrict__ TN) {
    using P = Plan<LOG2L>;
    constexpr int NB = P::L / R;                      // butterflies per half
    constexpr int TSTEP = 2 * (P::L / (S * R));       // index step into TN for W_(S*R)
    for (int bf = threadIdx.x; bf < NB; bf += kFftThreads) {
        const int low = bf & (S - 1), high = bf / S;
        const int base = high * S * R + low;
        float2 w = __ldg(&TN[low * TSTEP]);
        if (DIR > 0) w = cconj(w);
        float2 pw[R];
        twiddle_powers<R>(w, pw);
#pragma unroll
        for (int h = 0; h < 2; h++) {
            float2 *s = h ? sB : sA;
            float2 v[R];
#pragma unroll
            for (int k = 0; k < R; k++) v[k] = s[padidx(base + k * S)];
            if (DIR > 0) {
#pragma unroll
                for (int k = 1; k < R; k++) v[k] = cmul(v[k], pw[k]);
            }
            Dft<R, DIR>::run(v);
            if (DIR < 0) {
#pragma unroll
                for (int k = 1; k < R; k++) v[k] = cmul(v[k], pw[k]);
            }
#pragma unroll
            for (int k = 0; k < R; k++) s[padidx(base + k * S)] = v[k];
        }
    }
}

template <int LOG2L, int DIR>
__device__ __forceinline__ void mid_passes(float2 *sA, float2 *sB, const float2 *TN) {
    using P = Plan<LOG2L>;
    if (DIR < 0) {
        if constexpr (P::RA > 1) { mid_pass<LOG2L, P::RA, P::SA, -1>(sA, sB, TN); __syncthreads(); }
        if constexpr (P::RB > 1) { mid_pass<LOG2L, P::RB, 2, -1>(sA, sB, TN); __syncthreads(); }
    } else {
        if constexpr (P::RB > 1) { mid_pass<LOG2L, P::RB, 2, +1>(sA, sB, TN); __syncthreads(); }
        if constexpr (P::RA > 1) { mid_pass<LOG2L, P::RA, P::SA, +1>(sA, sB, TN); __syncthreads(); }
    }
}

// ---- K1: reference spectrum ------------------------------------------------------------------
// grid = blocks to transform; refspec layout [slot][2][L] float2 (even bins, odd bins), each in
// the digit-reversed order the forward passes produce.
template <int LOG2L>
__global__ void __launch_bounds__(kFftThreads, 1)
k_ref_spectrum(const uint8_t *__restrict__ iq, size_t block_stride, const uint32_t *__restrict__ block_list,
               uint32_t xormask, const float2 *__restrict__ TN, float2 *__restrict__ refspec) {
    using P = Plan<LOG2L>;
    extern __shared__ __align__(16) float2 smem[];
    float2 *sA = smem, *sB = smem + P::LP;
    const uint32_t slot = blockIdx.x;
    const uint32_t b = block_list ? block_list[slot] : slot;
    const uint8_t *x = iq + (size_t)b * block_stride;          // channel 0 of block b
    pass1_from_global<LOG2L, true>(x, xormask, sA, sB, TN);
    __syncthreads();
    mid_passes<LOG2L, -1>(sA, sB, TN);
    float2 *out = refspec + (size_t)slot * 2 * P::L;
    for (int m = threadIdx.x; m < P::L / 2; m += kFftThreads) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const float2 *s = h ? sB : sA;
            float4 pr = *reinterpret_cast<const float4 *>(&s[padidx(2 * m)]);
            float4 o = make_float4(pr.x + pr.z, pr.y + pr.w, pr.x - pr.z, pr.y - pr.w);
            *reinterpret_cast<float4 *>(&out[(size_t)h * P::L + 2 * m]) = o;
        }
    }
}

// ---- K2: correlate one (block, channel) item ---------------------------------------------------
template <int LOG2L>
__global__ void __launch_bounds__(kFftThreads, 1)
k_correlate(const uint8_t *__restrict__ iq, uint32_t nch, uint32_t B, const uint32_t *__restrict__ item_list,
            uint32_t n_signal, const uint32_t *__restrict__ block_slot, uint32_t xormask,
            const float2 *__restrict__ TN, const float2 *__restrict__ refspec, PeakRec *__restrict__ peaks) {
    using P = Plan<LOG2L>;
    constexpr int L = P::L, S = P::S1;
    extern __shared__ __align__(16) float2 smem[];
    __shared__ float s_val[kFftThreads / 32];
    __shared__ uint32_t s_idx[kFftThreads / 32];
    float2 *sA = smem, *sB = smem + P::LP;

    uint32_t b, c;
    if (item_list) { uint32_t it = item_list[blockIdx.x]; b = it / nch; c = it % nch; }
    else { b = blockIdx.x / n_signal; c = 1 + blockIdx.x % n_signal; }
    const uint8_t *x = iq + ((size_t)b * nch + c) * B;
    const uint32_t slot = block_slot ? block_slot[b] : b;
    const float2 *R = refspec + (size_t)slot * 2 * L;

    pass1_from_global<LOG2L, false>(x, xormask, sA, sB, TN);
    __syncthreads();
    mid_passes<LOG2L, -1>(sA, sB, TN);

    // last forward radix-2 stage, X * conj(R) (cdsp::conjugatemul, src/cdsp.cc:105-108),
    // first inverse radix-2 stage -- all on adjacent pairs, in registers
    for (int m = threadIdx.x; m < L / 2; m += kFftThreads) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            float2 *s = h ? sB : sA;
            float4 pr = *reinterpret_cast<const float4 *>(&s[padidx(2 * m)]);
            float4 rr = __ldg(reinterpret_cast<const float4 *>(&R[(size_t)h * L + 2 * m]));
            float2 y0 = make_float2(pr.x + pr.z, pr.y + pr.w), y1 = make_float2(pr.x - pr.z, pr.y - pr.w);
            float2 z0 = cmulc(y0, make_float2(rr.x, rr.y)), z1 = cmulc(y1, make_float2(rr.z, rr.w));
            float4 o = make_float4(z0.x + z1.x, z0.y + z1.y, z0.x - z1.x, z0.y - z1.y);
            *reinterpret_cast<float4 *>(&s[padidx(2 * m)]) = o;
        }
    }
    __syncthreads();
    mid_passes<LOG2L, +1>(sA, sB, TN);

    // last inverse pass (radix 16, stride L/16) kept in registers; recombine the halves,
    // |.|^2 (cdsp::magsquared, src/cdsp.cc:100-103) and running arg-max with lowest-index ties
    // (volk_32f_index_max_32u semantics, src/cdsp.cc:135-139)
    float best = -1.0f; uint32_t bidx = 0xffffffffu;
    for (int low = threadIdx.x; low < S; low += kFftThreads) {
        float2 a[16], bb[16];
        float2 wc = cconj(__ldg(&TN[2 * low])), gc = cconj(__ldg(&TN[low]));
        float2 pw[16];
        twiddle_powers<16>(wc, pw);
#pragma unroll
        for (int k = 0; k < 16; k++) {
            float2 va = sA[padidx(low + k * S)], vb = sB[padidx(low + k * S)];
            a[k] = (k == 0) ? va : cmul(va, pw[k]);
            bb[k] = cmul(vb, (k == 0) ? gc : cmul(gc, pw[k]));
        }
        Dft<16, +1>::run(a);
        Dft<16, +1>::run(bb);
#pragma unroll
        for (int k = 0; k < 16; k++) {
            float2 t = mul_w32<+1>(bb[k], k);
            float2 c0 = cadd(a[k], t), c1 = csub(a[k], t);
            float v0 = c0.x * c0.x + c0.y * c0.y, v1 = c1.x * c1.x + c1.y * c1.y;
            uint32_t i0 = low + k * S, i1 = i0 + L;
            if (v0 > best || (v0 == best && i0 < bidx)) { best = v0; bidx = i0; }
            if (v1 > best || (v1 == best && i1 < bidx)) { best = v1; bidx = i1; }
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        float ov = __shfl_down_sync(0xffffffffu, best, off);
        uint32_t oi = __shfl_down_sync(0xffffffffu, bidx, off);
        if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { s_val[warp] = best; s_idx[warp] = bidx; }
    __syncthreads();
    if (warp == 0) {
        best = lane < kFftThreads / 32 ? s_val[lane] : -1.0f;
        bidx = lane < kFftThreads / 32 ? s_idx[lane] : 0xffffffffu;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            float ov = __shfl_down_sync(0xffffffffu, best, off);
            uint32_t oi = __shfl_down_sync(0xffffffffu, bidx, off);
            if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
        }
        if (lane == 0) { PeakRec pr; pr.idx = bidx; pr.val = best; peaks[(size_t)b * nch + c] = pr; }
    }
}

// ---- host-side launchers ---------------------------------------------------------------------
template <int LOG2L>
static cudaError_t launch_ref_t(const CorrLaunch &p, cudaStream_t st) {
    using P = Plan<LOG2L>;
    static bool attr_done[64] = {};
    int dev = 0; cudaGetDevice(&dev);
    if (!attr_done[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(k_ref_spectrum<LOG2L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P::kSmemBytes);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_correlate<LOG2L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done[dev & 63] = true;
    }
    if (p.n_ref_blocks)
        k_ref_spectrum<LOG2L><<<p.n_ref_blocks, kFftThreads, P::kSmemBytes, st>>>(
            p.iq, (size_t)p.nch * p.B, p.ref_block_list, p.xormask, p.TN, p.refspec);
    if (p.n_items)
        k_correlate<LOG2L><<<p.n_items, kFftThreads, P::kSmemBytes, st>>>(
            p.iq, p.nch, p.B, p.item_list, p.nch - 1, p.block_slot, p.xormask, p.TN, p.refspec, p.peaks);
    return cudaGetLastError();
}

cudaError_t launch_correlate(const CorrLaunch &p, cudaStream_t st) {
    int log2L = 0;
    while ((1u << log2L) < p.B / 2) log2L++;
    switch (log2L) {
        case 5:  return launch_ref_t<5>(p, st);
        case 6:  return launch_ref_t<6>(p, st);
        case 7:  return launch_ref_t<7>(p, st);
        case 8:  return launch_ref_t<8>(p, st);
        case 9:  return launch_ref_t<9>(p, st);
        case 10: return launch_ref_t<10>(p, st);
        case 11: return launch_ref_t<11>(p, st);
        case 12: return launch_ref_t<12>(p, st);
        case 13: return launch_ref_t<13>(p, st);
        default: return cudaErrorInvalidValue;
    }
}

bool correlate_supported(uint32_t B) {
    uint32_t L = B / 2;
    return (L & (L - 1)) == 0 && L >= 32 && L <= 8192;
}

}  // namespace coh
