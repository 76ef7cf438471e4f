// fft_big.cu -- correlation for blocks whose transform does not fit one SM's shared memory
// (8192 < L <= 65536 complex samples per block; BASELINE config 3 is L = 65536, N = 2L = 131072).
//
// Same arithmetic as fft_kernels.cu (ccoherent::computelag, src/ccoherent.cc:174-192 of the
// reference: fft -> conjugatemul -> ifft -> magsquared -> indexofmax, on the zero-padded layout of
// src/crtlsdr.cc:205-218), split four-step style with L = Q * 8192, Q = 2, 4 or 8:
//   k_big_split : u8 -> cf32, first decimation-in-frequency stage of radix Q at stride 8192 and
//                 its twiddles; writes Q sub-sequences of 8192 packed elements to HBM scratch
//   k_big_sub   : one CTA per (item, sub-sequence): the 8192-point passes of fft_passes.cuh in
//                 shared memory -- forward, X * conj(R), inverse -- in place in the scratch
//                 (reference channel: forward only, into the spectrum ring)
//   k_big_merge : last inverse stage of radix Q, recombination of the two halves, |.|^2 and
//                 arg-max with lowest-index ties; one CTA per item
// The scratch round trips (4 x 16 L bytes per item) are the price of the size; they are counted
// as non-algorithmic traffic.
#include "fft_passes.cuh"

namespace coh {

constexpr int kSubLog2 = 13, kSub = 1 << kSubLog2;       // 8192-point sub-transforms
constexpr int kSplitThreads = 256;

struct BigAddr {                                         // which row of iq an item is
    const uint8_t *iq; uint32_t nch, B, n_signal;
    const uint32_t *item_list;                           // signal items: b*nch+c (NULL = all)
    const uint32_t *ref_block_list;                      // reference rows: block per slot (NULL = identity)
    const uint32_t *block_slot;                          // block -> spectrum slot (NULL = identity)
};
__device__ __forceinline__ void item_bc(const BigAddr &a, uint32_t it, uint32_t &b, uint32_t &c) {
    if (a.item_list) { const uint32_t v = a.item_list[it]; b = v / a.nch; c = v % a.nch; }
    else { b = it / a.n_signal; c = 1 + it % a.n_signal; }
}

// grid (8192 / 256, items); out[item][Q][8192]
template <int Q, bool IS_REF>
__global__ void __launch_bounds__(kSplitThreads)
k_big_split(const BigAddr a, uint32_t xormask, const float2 *__restrict__ TN, float4 *__restrict__ out) {
    const uint32_t it = blockIdx.y;
    uint32_t b, c;
    if (IS_REF) { b = a.ref_block_list ? a.ref_block_list[it] : it; c = 0; }
    else item_bc(a, it, b, c);
    const uint8_t *x = a.iq + ((size_t)b * a.nch + c) * a.B;
    const int low = blockIdx.x * kSplitThreads + threadIdx.x;
    PC v[Q];
#pragma unroll
    for (int k = 0; k < Q; k++) {
        const float2 s = load_sample(x, low + k * kSub, xormask);
        const float2 m = mul_w32<-1>(s, k * (16 / Q));    // W_N^(k*8192) = W_(2Q)^k
        v[k].r = make_float2(s.x, m.x); v[k].i = make_float2(s.y, m.y);
    }
    PDft<Q, -1>::run(v);
    float2 g = __ldg(&TN[low]);                           // W_N^low (odd-bin modulation)
    if (IS_REF) g = make_float2(-g.x, -g.y);              // ref sits in [L,2L): odd bins change sign
    const TwPair tw(__ldg(&TN[2 * low]), g);              // W_L^low
    float4 *o = out + (size_t)it * Q * kSub + low;
#pragma unroll
    for (int k = 0; k < Q; k++) o[(size_t)k * kSub] = pc_to(tw.apply(v[k], k));
}

// grid items * Q; data[item][Q][8192] transformed in place.  Same middle phase as the persistent L = 8192
// kernel (fft_kernels.cu): radix 16 at stride 32, fused 32-point stage with X * conj(R) and the first inverse
// 16-point DFT, inverse radix 16 at stride 32 -- five shared-memory round trips instead of seven; the reference
// spectrum of a sub-sequence is kept in the register order of that stage (fused_b_out).
constexpr size_t kBigSubSmem = ((size_t)ppadded_len_t<5>(kSub) * sizeof(float4) + 127) / 128 * 128;
template <bool REFONLY>
__global__ void __launch_bounds__(kFftThreads, 1)
k_big_sub(float4 *__restrict__ data, int Q, const BigAddr a, const float4 *__restrict__ refspec,
          const float2 *__restrict__ TNs) {
    using P = Plan<kSubLog2>;
    constexpr int S = P::S1;
    static_assert(S == kFftThreads, "one butterfly per thread in the first and last pass");
    extern __shared__ __align__(16) float4 smem4[];
    const uint32_t it = blockIdx.x / Q, q0 = blockIdx.x % Q;
    float4 *d = data + ((size_t)it * Q + q0) * kSub;
    const int low = threadIdx.x;
    const float2 w_s = __ldg(&TNs[2 * low]);                                    // W_8192^low
    const float2 w_a = mid_twiddle<kSubLog2, P::RA, P::SA>(TNs);                // W_512^(lane)
    const float4 *R = nullptr;
    if (!REFONLY) {
        uint32_t b, c;
        item_bc(a, it, b, c);
        const uint32_t slot = a.block_slot ? a.block_slot[b] : b;
        R = refspec + ((size_t)slot * Q + q0) * kSub;
    }
    {
        PC v[16];
#pragma unroll
        for (int k = 0; k < 16; k++) v[k] = pc_from(d[low + k * S]);
        PDft<16, -1>::run(v);
        const TwScalar tw(w_s);
#pragma unroll
        for (int k = 0; k < 16; k++) smem4[ppad_t<5>(low + k * S)] = pc_to(tw.apply(v[k], k));
    }
    __syncthreads();
    float4 r_early[kEarly ? kEarly : 1];
    if (!REFONLY) fused_b_early<kSubLog2>(R, r_early);
    mid_pass<kSubLog2, P::RA, P::SA, -1, 5>(smem4, w_a);
    __syncwarp();
    if (REFONLY) { fused_b_out<kSubLog2>(smem4, d); return; }
    fused_b_stage<kSubLog2>(smem4, R, r_early);
    __syncwarp();
    const int lane = threadIdx.x & 31;
    inv_a_stage<kSubLog2>(smem4, w_a, make_float2(kW32c[lane & 15], kW32s[lane & 15]));
    __syncthreads();
    {
        PC v[16];
        const TwScalar tw(cconj(w_s));
#pragma unroll
        for (int k = 0; k < 16; k++) v[k] = tw.apply(pc_from(smem4[ppad_t<5>(low + k * S)]), k);
        PDft<16, +1>::run(v);
#pragma unroll
        for (int k = 0; k < 16; k++) d[low + k * S] = pc_to(v[k]);
    }
}

// grid items; data[item][Q][8192] -> peaks[b*nch+c]
template <int Q>
__global__ void __launch_bounds__(kFftThreads)
k_big_merge(const float4 *__restrict__ data, const BigAddr a, const float2 *__restrict__ TN,
            PeakRec *__restrict__ peaks) {
    __shared__ float s_val[kFftThreads / 32];
    __shared__ uint32_t s_idx[kFftThreads / 32];
    const uint32_t it = blockIdx.x, L = Q * kSub;
    const float4 *d = data + (size_t)it * Q * kSub;
    float best = -1.0f; uint32_t bidx = 0xffffffffu;
    for (int low = threadIdx.x; low < kSub; low += kFftThreads) {
        PC v[Q];
        const TwPair tw(cconj(__ldg(&TN[2 * low])), cconj(__ldg(&TN[low])));
#pragma unroll
        for (int k = 0; k < Q; k++) v[k] = tw.apply(pc_from(__ldg(&d[(size_t)k * kSub + low])), k);
        PDft<Q, +1>::run(v);
#pragma unroll
        for (int k = 0; k < Q; k++) {
            // c[m] = A[m] + conj(W_N^m) B[m], c[m+L] = A[m] - conj(W_N^m) B[m], m = low + k*8192
            const float2 t = mul_w32<+1>(make_float2(v[k].r.y, v[k].i.y), k * (16 / Q));
            const float2 pm = make_float2(1.0f, -1.0f);
            const float2 cr = __ffma2_rn(splat(t.x), pm, splat(v[k].r.x));
            const float2 ci = __ffma2_rn(splat(t.y), pm, splat(v[k].i.x));
            const float2 mg = __ffma2_rn(ci, ci, __fmul2_rn(cr, cr));
            const uint32_t i0 = low + k * kSub, i1 = i0 + L;
            if (mg.x > best || (mg.x == best && i0 < bidx)) { best = mg.x; bidx = i0; }
            if (mg.y > best || (mg.y == best && i1 < bidx)) { best = mg.y; bidx = i1; }
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
        if (lane == 0) {
            uint32_t b, c;
            item_bc(a, it, b, c);
            PeakRec pr; pr.idx = bidx; pr.val = best; peaks[(size_t)b * a.nch + c] = pr;
        }
    }
}

// ---- host-side launcher ------------------------------------------------------------------------
template <int Q>
static cudaError_t launch_big_t(const CorrLaunch &p, cudaStream_t st) {
    using P = Plan<kSubLog2>;
    static bool attr_done[64] = {};
    int dev = 0; cudaGetDevice(&dev);
    if (!attr_done[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(k_big_sub<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBigSubSmem);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_big_sub<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBigSubSmem);
        if (e != cudaSuccess) return e;
        attr_done[dev & 63] = true;
    }
    BigAddr a{};
    a.iq = p.iq; a.nch = p.nch; a.B = p.B; a.n_signal = p.nch - 1;
    a.item_list = p.item_list; a.ref_block_list = p.ref_block_list; a.block_slot = p.block_slot;
    float4 *spec = reinterpret_cast<float4 *>(p.refspec);
    if (p.n_ref_blocks) {
        k_big_split<Q, true><<<dim3(kSub / kSplitThreads, p.n_ref_blocks), kSplitThreads, 0, st>>>(a, p.xormask, p.TN, spec);
        k_big_sub<true><<<p.n_ref_blocks * Q, kFftThreads, kBigSubSmem, st>>>(spec, Q, a, nullptr, p.TN_sub);
    }
    if (p.n_items) {
        if (!p.scratch || !p.TN_sub) return cudaErrorInvalidValue;
        k_big_split<Q, false><<<dim3(kSub / kSplitThreads, p.n_items), kSplitThreads, 0, st>>>(a, p.xormask, p.TN, p.scratch);
        k_big_sub<false><<<p.n_items * Q, kFftThreads, kBigSubSmem, st>>>(p.scratch, Q, a, spec, p.TN_sub);
        k_big_merge<Q><<<p.n_items, kFftThreads, 0, st>>>(p.scratch, a, p.TN, p.peaks);
    }
    return cudaGetLastError();
}

cudaError_t launch_correlate_big(const CorrLaunch &p, cudaStream_t st) {
    switch (p.B / 2 / kSub) {
        case 2: return launch_big_t<2>(p, st);
        case 4: return launch_big_t<4>(p, st);
        case 8: return launch_big_t<8>(p, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace coh
