// fft_r8.cu -- the L = 8192 correlation with radix-8 passes and 1024 threads per item.
//
// Same arithmetic and data layout as fft_kernels.cu (two packed halves of the zero-padded
// 2L-point transform, ccoherent::computelag src/ccoherent.cc:174-192), different schedule: the
// radix-16 kernel needs ~125 registers per thread, i.e. 16 warps per SM, and its FMA pipe idles
// while those warps are in their load/store phases.  Radix-8 butterflies (8 packed elements = 32
// registers) fit 64 registers, so 32 warps share the SM and one warp's shared-memory phase hides
// behind another's arithmetic.  8192 = 8*8*8*8*2:
//   pass 1   global -> shared, stride 1024         (block barrier)
//   mid A    stride 128, inside 1024-element sub-transforms = 128-thread groups (named barrier)
//   mid B    stride 16,  inside 128-element groups  = 16 threads (warp sync)
//   mid C    stride 2,   inside 16-element groups   = 2 threads  (warp sync)
//   pair     last forward radix-2, X * conj(R), first inverse radix-2 (warp sync)
//   mid C, B, A inverse, last pass stride 1024 in registers -> |.|^2 -> arg-max
// The reference spectrum must come out of the SAME pass structure (its order is the digit
// reversal of these radices), hence k_ref_spectrum_r8.
#include "fft_passes.cuh"

namespace coh {

constexpr int kT8 = 1024;                    // threads per item
constexpr int kL8 = 8192;
// two pad elements per 16: keeps the 8 lanes of every LDS.128 / STS.128 phase on distinct
// 16-byte bank groups for the strides 1024, 128, 16 and 2 of the radix-8 passes
__device__ __forceinline__ int pad8(int a) { return a + 2 * (a >> 4); }
constexpr int kLP8 = kL8 + 2 * (kL8 >> 4) + 2;
constexpr size_t kSmem8 = (size_t)kLP8 * sizeof(float4);

__device__ __forceinline__ void group_sync() {           // the 128 threads of one 1024-element sub-transform
    asm volatile("bar.sync %0, 128;" ::"r"(1 + (int)(threadIdx.x >> 7)) : "memory");
}

struct Tw8 { float2 w_l, w_n, a, b, c; };               // per-thread twiddle bases of all passes
__device__ __forceinline__ Tw8 load_tw8(const float2 *__restrict__ TN) {
    const int t = threadIdx.x;
    Tw8 w;
    w.w_l = __ldg(&TN[2 * t]);                           // W_L^low   (pass 1 / last pass)
    w.w_n = __ldg(&TN[t]);                               // W_N^low   (odd-bin modulation)
    w.a = __ldg(&TN[(t & 127) * 16]);                    // W_1024^low
    w.b = __ldg(&TN[(t & 15) * 128]);                    // W_128^low
    w.c = __ldg(&TN[(t & 1) * 1024]);                    // W_16^low
    return w;
}

template <bool NEG_B>
__device__ __forceinline__ void pass1_r8(const uint8_t *x, uint32_t xormask, float4 *s, float2 w_l, float2 w_n) {
    const int low = threadIdx.x;
    PC v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float2 a = load_sample(x, low + k * 1024, xormask);
        const float2 b = mul_w32<-1>(a, 2 * k);          // W_N^(k*1024) = W_16^k
        v[k].r = make_float2(a.x, b.x); v[k].i = make_float2(a.y, b.y);
    }
    PDft<8, -1>::run(v);
    float2 g = w_n;
    if (NEG_B) g = make_float2(-g.x, -g.y);              // ref sits in [L,2L): odd bins change sign
    const TwPair tw(w_l, g);
#pragma unroll
    for (int k = 0; k < 8; k++) s[pad8(low + k * 1024)] = pc_to(tw.apply(v[k], k));
}

template <int S, int DIR>
__device__ __forceinline__ void mid_r8(float4 *s, float2 w) {
    const int bf = threadIdx.x, low = bf & (S - 1), high = bf / S;
    const int base = high * S * 8 + low;
    if (DIR > 0) w = cconj(w);
    const TwScalar tw(w);
    PC v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = pc_from(s[pad8(base + k * S)]);
    if (DIR > 0) {
#pragma unroll
        for (int k = 1; k < 8; k++) v[k] = tw.apply(v[k], k);
    }
    PDft<8, DIR>::run(v);
    if (DIR < 0) {
#pragma unroll
        for (int k = 1; k < 8; k++) v[k] = tw.apply(v[k], k);
    }
#pragma unroll
    for (int k = 0; k < 8; k++) s[pad8(base + k * S)] = pc_to(v[k]);
}

// forward passes after pass 1 (callers put a block barrier in front)
__device__ __forceinline__ void forward_mid_r8(float4 *s, const Tw8 &w) {
    mid_r8<128, -1>(s, w.a); group_sync();
    mid_r8<16, -1>(s, w.b); __syncwarp();
    mid_r8<2, -1>(s, w.c); __syncwarp();
}
// pair m handled by this thread in trip j: the 128 pairs of the warp's own 256 elements
__device__ __forceinline__ int pair_r8(int j) { return (threadIdx.x >> 5) * 128 + (threadIdx.x & 31) + 32 * j; }

// grid = blocks to transform; refspec [slot][L] float4 in the order forward_mid_r8 leaves
__global__ void __launch_bounds__(kT8, 1)
k_ref_spectrum_r8(const uint8_t *__restrict__ iq, size_t block_stride, const uint32_t *__restrict__ block_list,
                  uint32_t xormask, const float2 *__restrict__ TN, float4 *__restrict__ refspec) {
    extern __shared__ __align__(16) float4 smem4[];
    const uint32_t slot = blockIdx.x;
    const uint32_t b = block_list ? block_list[slot] : slot;
    const Tw8 w = load_tw8(TN);
    pass1_r8<true>(iq + (size_t)b * block_stride, xormask, smem4, w.w_l, w.w_n);
    __syncthreads();
    forward_mid_r8(smem4, w);
    float4 *out = refspec + (size_t)slot * kL8;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int m = pair_r8(j);
        const int sw = (threadIdx.x >> 2) & 1;           // see k_correlate_r8
        const float4 q0 = smem4[pad8(2 * m + sw)], q1 = smem4[pad8(2 * m + 1 - sw)];
        const PC e0 = pc_from(sw ? q1 : q0), e1 = pc_from(sw ? q0 : q1);
        out[2 * m] = pc_to(pc_add(e0, e1));
        out[2 * m + 1] = pc_to(pc_sub(e0, e1));
    }
}

// grid = items
__global__ void __launch_bounds__(kT8, 1)
k_correlate_r8(const uint8_t *__restrict__ iq, uint32_t nch, uint32_t B, const uint32_t *__restrict__ item_list,
               uint32_t n_signal, const uint32_t *__restrict__ block_slot, uint32_t xormask,
               const float2 *__restrict__ TN, const float4 *__restrict__ refspec, PeakRec *__restrict__ peaks) {
    constexpr int L = kL8;
    extern __shared__ __align__(16) float4 smem4[];
    __shared__ float s_val[kT8 / 32];
    __shared__ uint32_t s_idx[kT8 / 32];

    uint32_t b, c;
    if (item_list) { uint32_t it = item_list[blockIdx.x]; b = it / nch; c = it % nch; }
    else { b = blockIdx.x / n_signal; c = 1 + blockIdx.x % n_signal; }
    const uint8_t *x = iq + ((size_t)b * nch + c) * B;
    const uint32_t slot = block_slot ? block_slot[b] : b;
    const float4 *R = refspec + (size_t)slot * L;
    const Tw8 w = load_tw8(TN);

    pass1_r8<false>(x, xormask, smem4, w.w_l, w.w_n);
    __syncthreads();
    forward_mid_r8(smem4, w);
    {
        // last forward radix-2 stage, X * conj(R) (cdsp::conjugatemul, src/cdsp.cc:105-108), first
        // inverse radix-2 stage on adjacent pairs; all four spectrum rows in flight
        float4 g0[4], g1[4];
#pragma unroll
        for (int j = 0; j < 4; j++) { const int m = pair_r8(j); g0[j] = __ldg(&R[2 * m]); g1[j] = __ldg(&R[2 * m + 1]); }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int m = pair_r8(j);
            // lanes 0-3 of every quarter-warp touch the even element first, lanes 4-7 the odd one:
            // the eight 16-byte words of one LDS.128 / STS.128 phase fall on distinct bank groups
            const int sw = (threadIdx.x >> 2) & 1;
            const float4 q0 = smem4[pad8(2 * m + sw)], q1 = smem4[pad8(2 * m + 1 - sw)];
            const PC e0 = pc_from(sw ? q1 : q0), e1 = pc_from(sw ? q0 : q1);
            const PC z0 = pc_mulcp(pc_add(e0, e1), pc_from(g0[j])), z1 = pc_mulcp(pc_sub(e0, e1), pc_from(g1[j]));
            const float4 o0 = pc_to(pc_add(z0, z1)), o1 = pc_to(pc_sub(z0, z1));
            smem4[pad8(2 * m + sw)] = sw ? o1 : o0;
            smem4[pad8(2 * m + 1 - sw)] = sw ? o0 : o1;
        }
    }
    __syncwarp();
    mid_r8<2, +1>(smem4, w.c); __syncwarp();
    mid_r8<16, +1>(smem4, w.b); group_sync();
    mid_r8<128, +1>(smem4, w.a); __syncthreads();

    // last inverse pass (radix 8, stride 1024) in registers; recombine the halves, |.|^2
    // (cdsp::magsquared, src/cdsp.cc:100-103) and arg-max with lowest-index ties
    // (volk_32f_index_max_32u semantics, src/cdsp.cc:135-139)
    const int low = threadIdx.x;
    PC v[8];
    {
        const TwPair tw(cconj(w.w_l), cconj(w.w_n));
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = tw.apply(pc_from(smem4[pad8(low + k * 1024)]), k);
    }
    PDft<8, +1>::run(v);
    float2 mg[8];                                        // (|c[m]|^2, |c[m+L]|^2), m = low + k*1024
#pragma unroll
    for (int k = 0; k < 8; k++) {
        // c[m] = A[m] + conj(W_N^(k*1024)) B'[m], c[m+L] = A[m] - ...  (B' already carries conj(W_N^low))
        const float2 t = mul_w32<+1>(make_float2(v[k].r.y, v[k].i.y), 2 * k);
        const float2 pm = make_float2(1.0f, -1.0f);
        const float2 cr = __ffma2_rn(splat(t.x), pm, splat(v[k].r.x));
        const float2 ci = __ffma2_rn(splat(t.y), pm, splat(v[k].i.x));
        mg[k] = __ffma2_rn(ci, ci, __fmul2_rn(cr, cr));
    }
    float best = -1.0f; uint32_t bidx = 0xffffffffu;
    // indices visited in increasing order within the thread: strict > keeps the lowest
#pragma unroll
    for (int k = 0; k < 8; k++) if (mg[k].x > best) { best = mg[k].x; bidx = low + k * 1024; }
#pragma unroll
    for (int k = 0; k < 8; k++) if (mg[k].y > best) { best = mg[k].y; bidx = low + k * 1024 + L; }
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
        best = s_val[lane]; bidx = s_idx[lane];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            float ov = __shfl_down_sync(0xffffffffu, best, off);
            uint32_t oi = __shfl_down_sync(0xffffffffu, bidx, off);
            if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
        }
        if (lane == 0) { PeakRec pr; pr.idx = bidx; pr.val = best; peaks[(size_t)b * nch + c] = pr; }
    }
}

cudaError_t launch_correlate_r8(const CorrLaunch &p, cudaStream_t st) {
    static bool attr_done[64] = {};
    int dev = 0; cudaGetDevice(&dev);
    if (!attr_done[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(k_ref_spectrum_r8, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem8);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_correlate_r8, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem8);
        if (e != cudaSuccess) return e;
        attr_done[dev & 63] = true;
    }
    if (p.n_ref_blocks)
        k_ref_spectrum_r8<<<p.n_ref_blocks, kT8, kSmem8, st>>>(
            p.iq, (size_t)p.nch * p.B, p.ref_block_list, p.xormask, p.TN, reinterpret_cast<float4 *>(p.refspec));
    if (p.n_items)
        k_correlate_r8<<<p.n_items, kT8, kSmem8, st>>>(
            p.iq, p.nch, p.B, p.item_list, p.nch - 1, p.block_slot, p.xormask, p.TN,
            reinterpret_cast<const float4 *>(p.refspec), p.peaks);
    return cudaGetLastError();
}

}  // namespace coh
