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
// in shared memory -- one float4 {A.re, B.re, A.im, B.im} per element -- and every pass processes
// both at once on the packed fp32 pipe (fft_packed.cuh): one FADD2 / FFMA2 serves both halves.
// Forward passes are decimation-in-frequency (natural in, digit-reversed out), inverse passes
// decimation-in-time (digit-reversed in, natural out), so no reordering pass exists and the
// reference spectrum is stored in the same digit-reversed order.
#include "fft_passes.cuh"

namespace coh {

// ---- K1: reference spectrum ------------------------------------------------------------------
// grid = blocks to transform; refspec layout [slot][L] float4 {even.re, odd.re, even.im, odd.im}
// in the digit-reversed order the forward passes produce (the order k_correlate consumes).
template <int LOG2L>
__global__ void __launch_bounds__(kFftThreads, 1)
k_ref_spectrum(const uint8_t *__restrict__ iq, size_t block_stride, const uint32_t *__restrict__ block_list,
               uint32_t xormask, const float2 *__restrict__ TN, float4 *__restrict__ refspec) {
    using P = Plan<LOG2L>;
    extern __shared__ __align__(16) float4 smem4[];
    const uint32_t slot = blockIdx.x;
    const uint32_t b = block_list ? block_list[slot] : slot;
    const uint8_t *x = iq + (size_t)b * block_stride;          // channel 0 of block b
    pass1_from_global<LOG2L, true>(x, xormask, smem4, TN);
    __syncthreads();
    mid_passes<LOG2L, -1>(smem4, TN);
    pair_out<LOG2L>(smem4, refspec + (size_t)slot * P::L);
}

// ---- K2: correlate one (block, channel) item ---------------------------------------------------
template <int LOG2L>
__global__ void __launch_bounds__(kFftThreads, 1)
k_correlate(const uint8_t *__restrict__ iq, uint32_t nch, uint32_t B, const uint32_t *__restrict__ item_list,
            uint32_t n_signal, const uint32_t *__restrict__ block_slot, uint32_t xormask,
            const float2 *__restrict__ TN, const float4 *__restrict__ refspec, PeakRec *__restrict__ peaks) {
    using P = Plan<LOG2L>;
    constexpr int L = P::L, S = P::S1;
    extern __shared__ __align__(16) float4 smem4[];
    __shared__ float s_val[kFftThreads / 32];
    __shared__ uint32_t s_idx[kFftThreads / 32];

    uint32_t b, c;
    if (item_list) { uint32_t it = item_list[blockIdx.x]; b = it / nch; c = it % nch; }
    else { b = blockIdx.x / n_signal; c = 1 + blockIdx.x % n_signal; }
    const uint8_t *x = iq + ((size_t)b * nch + c) * B;
    const uint32_t slot = block_slot ? block_slot[b] : b;
    const float4 *R = refspec + (size_t)slot * L;

    // per-thread twiddle bases of all passes, fetched once ahead of the barriers
    static_assert(S <= kFftThreads, "one butterfly per thread in every pass");
    const int tlow = threadIdx.x < S ? threadIdx.x : 0;
    const float2 w_l = __ldg(&TN[2 * tlow]), w_n = __ldg(&TN[tlow]);   // W_L^low, W_N^low
    const MidTw mt = mid_twiddles<LOG2L>(TN);

    pass1_from_global<LOG2L, false>(x, xormask, smem4, w_l, w_n);
    __syncthreads();
    mid_passes<LOG2L, -1>(smem4, mt);

    pair_correlate<LOG2L>(smem4, R);
    mid_sync<LOG2L>();
    mid_passes<LOG2L, +1>(smem4, mt);
    if constexpr (WarpLocal<LOG2L>::value) __syncthreads();

    // last inverse pass (radix 16, stride L/16) kept in registers; recombine the halves,
    // |.|^2 (cdsp::magsquared, src/cdsp.cc:100-103) and running arg-max with lowest-index ties
    // (volk_32f_index_max_32u semantics, src/cdsp.cc:135-139)
    float best = -1.0f; uint32_t bidx = 0xffffffffu;
    for (int low = threadIdx.x; low < S; low += kFftThreads) {
        PC v[16];
        {
            const TwPair tw(cconj(w_l), cconj(w_n));
#pragma unroll
            for (int k = 0; k < 16; k++) v[k] = tw.apply(pc_from(smem4[ppad(low + k * S)]), k);
        }
        PDft<16, +1>::run(v);
        float2 mg[16];                                 // (|c[m]|^2, |c[m+L]|^2), m = low + k*S
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const float2 t = mul_w32<+1>(make_float2(v[k].r.y, v[k].i.y), k);
            const float2 pm = make_float2(1.0f, -1.0f);
            const float2 cr = __ffma2_rn(splat(t.x), pm, splat(v[k].r.x));
            const float2 ci = __ffma2_rn(splat(t.y), pm, splat(v[k].i.x));
            mg[k] = __ffma2_rn(ci, ci, __fmul2_rn(cr, cr));
        }
        // indices visited in increasing order within the thread: strict > keeps the lowest
        if (S <= kFftThreads) {
#pragma unroll
            for (int k = 0; k < 16; k++) if (mg[k].x > best) { best = mg[k].x; bidx = low + k * S; }
#pragma unroll
            for (int k = 0; k < 16; k++) if (mg[k].y > best) { best = mg[k].y; bidx = low + k * S + L; }
        } else {
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const uint32_t i0 = low + k * S, i1 = i0 + L;
                if (mg[k].x > best || (mg[k].x == best && i0 < bidx)) { best = mg[k].x; bidx = i0; }
                if (mg[k].y > best || (mg[k].y == best && i1 < bidx)) { best = mg[k].y; bidx = i1; }
            }
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
            p.iq, (size_t)p.nch * p.B, p.ref_block_list, p.xormask, p.TN, reinterpret_cast<float4 *>(p.refspec));
    if (p.n_items)
        k_correlate<LOG2L><<<p.n_items, kFftThreads, P::kSmemBytes, st>>>(
            p.iq, p.nch, p.B, p.item_list, p.nch - 1, p.block_slot, p.xormask, p.TN, reinterpret_cast<const float4 *>(p.refspec), p.peaks);
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
        case 14: case 15: case 16: return launch_correlate_big(p, st);
        default: return cudaErrorInvalidValue;
    }
}

bool correlate_supported(uint32_t B) {
    uint32_t L = B / 2;
    return (L & (L - 1)) == 0 && L >= 32 && L <= 65536;
}
bool correlate_is_big(uint32_t B) { return B / 2 > 8192; }

}  // namespace coh
