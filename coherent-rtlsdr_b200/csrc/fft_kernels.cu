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
#include "fft_dev.cuh"
#include "fft_packed.cuh"
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
    static constexpr int LP = ppadded_len(L);
    static constexpr size_t kSmemBytes = (size_t)LP * sizeof(float4);
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

// Twiddles of one butterfly, two-level so that few registers stay live: w^(4a+b) = w^(4a) * w^b.
// Scalar form (the same twiddle for both halves) ...
struct TwScalar {
    float2 lo[4], hi[4];                            // w^0..w^3, w^0, w^4, w^8, w^12
    __device__ __forceinline__ explicit TwScalar(float2 w) {
        lo[0] = hi[0] = make_float2(1.0f, 0.0f);
        lo[1] = w; lo[2] = cmul(w, w); lo[3] = cmul(lo[2], w);
        hi[1] = cmul(lo[2], lo[2]); hi[2] = cmul(hi[1], hi[1]); hi[3] = cmul(hi[2], hi[1]);
    }
    __device__ __forceinline__ PC apply(PC v, int k) const {
        if (k == 0) return v;
        if (k < 4) return pc_mul(v, lo[k]);
        if ((k & 3) == 0) return pc_mul(v, hi[k >> 2]);
        return pc_mul(v, cmul(hi[k >> 2], lo[k & 3]));
    }
};
// ... and packed form: lane A gets w^k, lane B gets g * w^k (the odd-bin modulation rides along)
struct TwPair {
    PC lo[4]; float2 hi[4];
    __device__ __forceinline__ TwPair(float2 w, float2 g) {
        const float2 w2 = cmul(w, w), w3 = cmul(w2, w);
        lo[0].r = make_float2(1.0f, g.x); lo[0].i = make_float2(0.0f, g.y);
        lo[1] = pc_mul(lo[0], w); lo[2] = pc_mul(lo[0], w2); lo[3] = pc_mul(lo[0], w3);
        hi[0] = make_float2(1.0f, 0.0f);
        hi[1] = cmul(w2, w2); hi[2] = cmul(hi[1], hi[1]); hi[3] = cmul(hi[2], hi[1]);
    }
    __device__ __forceinline__ PC apply(PC v, int k) const {
        if (k < 4) return pc_mulp(v, lo[k]);
        return pc_mulp(v, pc_mul(lo[k & 3], hi[k >> 2]));
    }
};

// ---- pass 1: global bytes -> shared, radix 16 at stride L/16, both halves ------------------
template <int LOG2L, bool NEG_B>
__device__ __forceinline__ void pass1_from_global(const uint8_t *x, uint32_t xormask, float4 *s,
                                                  const float2 *__restrict__ TN) {
    using P = Plan<LOG2L>;
    constexpr int S = P::S1;
    for (int low = threadIdx.x; low < S; low += kFftThreads) {
        PC v[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const float2 a = load_sample(x, low + k * S, xormask);
            const float2 b = mul_w32<-1>(a, k);       // W_N^(k*L/16) = W_32^k
            v[k].r = make_float2(a.x, b.x); v[k].i = make_float2(a.y, b.y);
        }
        PDft<16, -1>::run(v);
        float2 g = __ldg(&TN[low]);                   // W_N^low (odd-bin modulation)
        if (NEG_B) g = make_float2(-g.x, -g.y);       // ref sits in [L,2L): odd bins change sign
        const TwPair tw(__ldg(&TN[2 * low]), g);      // W_L^low
#pragma unroll
        for (int k = 0; k < 16; k++) s[ppad(low + k * S)] = pc_to(tw.apply(v[k], k));
    }
}

// ---- in-place mid passes ----------------------------------------------------------------------
template <int LOG2L, int R, int S, int DIR>
__device__ __forceinline__ void mid_pass(float4 *s, const float2 *__restrict__ TN) {
    using P = Plan<LOG2L>;
    constexpr int NB = P::L / R;                      // butterflies
    constexpr int TSTEP = 2 * (P::L / (S * R));       // index step into TN for W_(S*R)
    for (int bf = threadIdx.x; bf < NB; bf += kFftThreads) {
        const int low = bf & (S - 1), high = bf / S;
        const int base = high * S * R + low;
        float2 w = __ldg(&TN[low * TSTEP]);
        if (DIR > 0) w = cconj(w);
        const TwScalar tw(w);
        PC v[R];
#pragma unroll
        for (int k = 0; k < R; k++) v[k] = pc_from(s[ppad(base + k * S)]);
        if (DIR > 0) {
#pragma unroll
            for (int k = 1; k < R; k++) v[k] = tw.apply(v[k], k);
        }
        PDft<R, DIR>::run(v);
        if (DIR < 0) {
#pragma unroll
            for (int k = 1; k < R; k++) v[k] = tw.apply(v[k], k);
        }
#pragma unroll
        for (int k = 0; k < R; k++) s[ppad(base + k * S)] = pc_to(v[k]);
    }
}

// After pass 1 the data is 16 independent sub-transforms of L/16 points; for L = 8192 with 512
// threads each of them (512 elements) is worked on by exactly one warp in every mid pass and
// in the pair stage (butterfly bf = thread index), so those passes need only warp-level
// synchronisation.  The 16 warps then drift apart and one warp's LDS/STS bursts overlap
// another's arithmetic instead of all warps meeting at a block barrier seven times per item.
template <int LOG2L> struct WarpLocal { static constexpr bool value = (LOG2L == 13) && (kFftThreads == 512); };
template <int LOG2L> __device__ __forceinline__ void mid_sync() {
    if constexpr (WarpLocal<LOG2L>::value) __syncwarp(); else __syncthreads();
}
// pair index handled by this thread in trip j of the pair stage
template <int LOG2L> __device__ __forceinline__ int pair_index(int j) {
    if constexpr (WarpLocal<LOG2L>::value) return (threadIdx.x >> 5) * 256 + (threadIdx.x & 31) + 32 * j;
    else return threadIdx.x + kFftThreads * j;
}

template <int LOG2L, int DIR>
__device__ __forceinline__ void mid_passes(float4 *s, const float2 *TN) {
    using P = Plan<LOG2L>;
    if (DIR < 0) {
        if constexpr (P::RA > 1) { mid_pass<LOG2L, P::RA, P::SA, -1>(s, TN); mid_sync<LOG2L>(); }
        if constexpr (P::RB > 1) { mid_pass<LOG2L, P::RB, 2, -1>(s, TN); mid_sync<LOG2L>(); }
    } else {
        if constexpr (P::RB > 1) { mid_pass<LOG2L, P::RB, 2, +1>(s, TN); mid_sync<LOG2L>(); }
        if constexpr (P::RA > 1) { mid_pass<LOG2L, P::RA, P::SA, +1>(s, TN); mid_sync<LOG2L>(); }
    }
}

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
    float4 *out = refspec + (size_t)slot * P::L;
    constexpr int kTrips = (P::L / 2 + kFftThreads - 1) / kFftThreads;
    for (int j = 0; j < kTrips; j++) {
        const int m = pair_index<LOG2L>(j);
        if (m >= P::L / 2) break;
        const PC e0 = pc_from(smem4[ppad(2 * m)]), e1 = pc_from(smem4[ppad(2 * m + 1)]);
        out[2 * m] = pc_to(pc_add(e0, e1));
        out[2 * m + 1] = pc_to(pc_sub(e0, e1));
    }
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

    pass1_from_global<LOG2L, false>(x, xormask, smem4, TN);
    __syncthreads();
    mid_passes<LOG2L, -1>(smem4, TN);

    // last forward radix-2 stage, X * conj(R) (cdsp::conjugatemul, src/cdsp.cc:105-108),
    // first inverse radix-2 stage -- all on adjacent pairs, in registers
    constexpr int kTrips = (L / 2 + kFftThreads - 1) / kFftThreads;
    for (int j = 0; j < kTrips; j++) {
        const int m = pair_index<LOG2L>(j);
        if (m >= L / 2) break;
        // lanes 0-3 of every quarter-warp touch the even element first, lanes 4-7 the odd one:
        // the eight 16-byte words of one LDS.128 / STS.128 phase then fall on distinct bank groups
        const int sw = (threadIdx.x >> 2) & 1;
        const float4 q0 = smem4[ppad(2 * m + sw)], q1 = smem4[ppad(2 * m + 1 - sw)];
        const PC e0 = pc_from(sw ? q1 : q0), e1 = pc_from(sw ? q0 : q1);
        const PC r0 = pc_from(__ldg(&R[2 * m])), r1 = pc_from(__ldg(&R[2 * m + 1]));
        const PC z0 = pc_mulcp(pc_add(e0, e1), r0), z1 = pc_mulcp(pc_sub(e0, e1), r1);
        const float4 o0 = pc_to(pc_add(z0, z1)), o1 = pc_to(pc_sub(z0, z1));
        smem4[ppad(2 * m + sw)] = sw ? o1 : o0;
        smem4[ppad(2 * m + 1 - sw)] = sw ? o0 : o1;
    }
    mid_sync<LOG2L>();
    mid_passes<LOG2L, +1>(smem4, TN);
    if constexpr (WarpLocal<LOG2L>::value) __syncthreads();

    // last inverse pass (radix 16, stride L/16) kept in registers; recombine the halves,
    // |.|^2 (cdsp::magsquared, src/cdsp.cc:100-103) and running arg-max with lowest-index ties
    // (volk_32f_index_max_32u semantics, src/cdsp.cc:135-139)
    float best = -1.0f; uint32_t bidx = 0xffffffffu;
    for (int low = threadIdx.x; low < S; low += kFftThreads) {
        PC v[16];
        {
            const TwPair tw(cconj(__ldg(&TN[2 * low])), cconj(__ldg(&TN[low])));
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
        default: return cudaErrorInvalidValue;
    }
}

bool correlate_supported(uint32_t B) {
    uint32_t L = B / 2;
    return (L & (L - 1)) == 0 && L >= 32 && L <= 8192;
}

}  // namespace coh
