// fft_p2.cu -- the L = 8192 correlation, radix 16, as TWO interleaved instruction streams per thread.
//
// ncu on k_correlate<13> (profiles/r1d_ncu_k_correlate.txt): FMA pipe ~55 % busy, shared-memory pipe
// ~45 % busy -- they add up to one: the 16 warps meet at the same barriers, so their load/store
// phases and their butterfly phases line up instead of overlapping.  Here each thread owns two
// butterflies per pass that belong to two INDEPENDENT halves of the data (after pass 1 the item is
// 16 separate 512-point sub-transforms; stream X works on sub-transforms 0-7, stream Y on 8-15, each
// sub-transform stays inside one warp).  The code alternates: while X's butterflies run, Y's
// LDS.128s are in flight, and the other way round, so every warp overlaps its own memory and
// arithmetic phases by instruction-level parallelism: 256 threads, up to 255 registers.
// Same passes, same twiddles, same spectrum order as fft_kernels.cu (k_ref_spectrum<13> feeds it).
#include "fft_passes.cuh"

namespace coh {

constexpr int kT2 = 256;
using P13 = Plan<13>;

template <int S>
__device__ __forceinline__ void p2_load(PC (&v)[16], const float4 *s, int base) {
#pragma unroll
    for (int k = 0; k < 16; k++) v[k] = pc_from(s[ppad(base + k * S)]);
}
template <int S>
__device__ __forceinline__ void p2_store(const PC (&v)[16], float4 *s, int base) {
#pragma unroll
    for (int k = 0; k < 16; k++) s[ppad(base + k * S)] = pc_to(v[k]);
}
__device__ __forceinline__ void p2_fwd(PC (&v)[16], const TwScalar &tw) {
    PDft<16, -1>::run(v);
#pragma unroll
    for (int k = 1; k < 16; k++) v[k] = tw.apply(v[k], k);
}
__device__ __forceinline__ void p2_inv(PC (&v)[16], const TwScalar &tw) {
#pragma unroll
    for (int k = 1; k < 16; k++) v[k] = tw.apply(v[k], k);
    PDft<16, +1>::run(v);
}

// pair stage of one stream: the 256 pairs of each of the warp's sub-transforms, four trips of
// spectrum loads in flight (same lane swizzle as pair_correlate)
__device__ __forceinline__ void p2_pair(float4 *s, const float4 *__restrict__ R, int seg) {
    const int lane = threadIdx.x & 31, sw = (threadIdx.x >> 2) & 1;
#pragma unroll 1
    for (int j0 = 0; j0 < 8; j0 += 4) {
        float4 g0[4], g1[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int m = seg * 256 + lane + 32 * (j0 + u);
            g0[u] = __ldg(&R[2 * m]); g1[u] = __ldg(&R[2 * m + 1]);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int m = seg * 256 + lane + 32 * (j0 + u);
            const float4 q0 = s[ppad(2 * m + sw)], q1 = s[ppad(2 * m + 1 - sw)];
            const PC e0 = pc_from(sw ? q1 : q0), e1 = pc_from(sw ? q0 : q1);
            const PC z0 = pc_mulcp(pc_add(e0, e1), pc_from(g0[u])), z1 = pc_mulcp(pc_sub(e0, e1), pc_from(g1[u]));
            const float4 o0 = pc_to(pc_add(z0, z1)), o1 = pc_to(pc_sub(z0, z1));
            s[ppad(2 * m + sw)] = sw ? o1 : o0;
            s[ppad(2 * m + 1 - sw)] = sw ? o0 : o1;
        }
    }
}

__device__ __forceinline__ void p2_peak(const PC (&v)[16], int low, float &best, uint32_t &bidx) {
    constexpr int S = P13::S1, L = P13::L;
    float2 mg[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const float2 t = mul_w32<+1>(make_float2(v[k].r.y, v[k].i.y), k);
        const float2 pm = make_float2(1.0f, -1.0f);
        const float2 cr = __ffma2_rn(splat(t.x), pm, splat(v[k].r.x));
        const float2 ci = __ffma2_rn(splat(t.y), pm, splat(v[k].i.x));
        mg[k] = __ffma2_rn(ci, ci, __fmul2_rn(cr, cr));
    }
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const uint32_t i0 = low + k * S, i1 = i0 + L;
        if (mg[k].x > best || (mg[k].x == best && i0 < bidx)) { best = mg[k].x; bidx = i0; }
        if (mg[k].y > best || (mg[k].y == best && i1 < bidx)) { best = mg[k].y; bidx = i1; }
    }
}

__global__ void __launch_bounds__(kT2, 1)
k_correlate_p2(const uint8_t *__restrict__ iq, uint32_t nch, uint32_t B, const uint32_t *__restrict__ item_list,
               uint32_t n_signal, const uint32_t *__restrict__ block_slot, uint32_t xormask,
               const float2 *__restrict__ TN, const float4 *__restrict__ refspec, PeakRec *__restrict__ peaks) {
    constexpr int L = P13::L, S = P13::S1;
    extern __shared__ __align__(16) float4 smem4[];
    __shared__ float s_val[kT2 / 32];
    __shared__ uint32_t s_idx[kT2 / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    uint32_t b, c;
    if (item_list) { uint32_t it = item_list[blockIdx.x]; b = it / nch; c = it % nch; }
    else { b = blockIdx.x / n_signal; c = 1 + blockIdx.x % n_signal; }
    const uint8_t *x = iq + ((size_t)b * nch + c) * B;
    const uint32_t slot = block_slot ? block_slot[b] : b;
    const float4 *R = refspec + (size_t)slot * L;

    // butterflies of the two streams in every pass: bf = tid and tid + 256
    const int lowX = tid, lowY = tid + kT2;
    const float2 wA = __ldg(&TN[lane * 32]);              // mid A: W_512^lane, shared by both streams
    const float2 wB = __ldg(&TN[(tid & 1) * 512]);        // mid B: W_32^(tid & 1)
    const float2 wlX = __ldg(&TN[2 * lowX]), wnX = __ldg(&TN[lowX]);
    const float2 wlY = __ldg(&TN[2 * lowY]), wnY = __ldg(&TN[lowY]);
    pass1_butterfly<13, false>(x, xormask, smem4, lowX, wlX, wnX);
    pass1_butterfly<13, false>(x, xormask, smem4, lowY, wlY, wnY);
    __syncthreads();

    // sub-transform (512 elements) of each stream and the butterfly bases inside it
    const int segX = warp, segY = warp + 8;
    const int aX = segX * 512 + lane, aY = segY * 512 + lane;                     // mid A: + 32 k
    const int bX = (lowX >> 1) * 32 + (tid & 1), bY = (lowY >> 1) * 32 + (tid & 1);   // mid B: + 2 k
    PC vx[16], vy[16];
    {
        const TwScalar tA(wA);
        p2_load<32>(vx, smem4, aX);
        p2_load<32>(vy, smem4, aY);
        p2_fwd(vx, tA); p2_store<32>(vx, smem4, aX);
        __syncwarp();
        p2_load<2>(vx, smem4, bX);                       // X's next loads fly while Y computes
        p2_fwd(vy, tA); p2_store<32>(vy, smem4, aY);
        __syncwarp();
    }
    {
        const TwScalar tB(wB);
        p2_load<2>(vy, smem4, bY);
        p2_fwd(vx, tB); p2_store<2>(vx, smem4, bX);
        __syncwarp();
        p2_fwd(vy, tB); p2_store<2>(vy, smem4, bY);
        p2_pair(smem4, R, segX);                         // X's pair stage overlaps the tail of Y's butterflies
        __syncwarp();
        p2_load<2>(vx, smem4, bX);
        p2_pair(smem4, R, segY);
        __syncwarp();
        const TwScalar tBi(cconj(wB));
        p2_load<2>(vy, smem4, bY);
        p2_inv(vx, tBi); p2_store<2>(vx, smem4, bX);
        __syncwarp();
        p2_load<32>(vx, smem4, aX);
        p2_inv(vy, tBi); p2_store<2>(vy, smem4, bY);
        __syncwarp();
    }
    {
        const TwScalar tAi(cconj(wA));
        p2_load<32>(vy, smem4, aY);
        p2_inv(vx, tAi); p2_store<32>(vx, smem4, aX);
        p2_inv(vy, tAi); p2_store<32>(vy, smem4, aY);
    }
    __syncthreads();

    // last inverse pass (radix 16, stride 512) in registers, both streams; |.|^2 and arg-max
    float best = -1.0f; uint32_t bidx = 0xffffffffu;
    {
        const TwPair twX(cconj(wlX), cconj(wnX));
#pragma unroll
        for (int k = 0; k < 16; k++) vx[k] = twX.apply(pc_from(smem4[ppad(lowX + k * S)]), k);
    }
    {
        const TwPair twY(cconj(wlY), cconj(wnY));
#pragma unroll
        for (int k = 0; k < 16; k++) vy[k] = twY.apply(pc_from(smem4[ppad(lowY + k * S)]), k);
    }
    PDft<16, +1>::run(vx);
    p2_peak(vx, lowX, best, bidx);
    PDft<16, +1>::run(vy);
    p2_peak(vy, lowY, best, bidx);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        float ov = __shfl_down_sync(0xffffffffu, best, off);
        uint32_t oi = __shfl_down_sync(0xffffffffu, bidx, off);
        if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
    }
    if (lane == 0) { s_val[warp] = best; s_idx[warp] = bidx; }
    __syncthreads();
    if (warp == 0) {
        best = lane < kT2 / 32 ? s_val[lane] : -1.0f;
        bidx = lane < kT2 / 32 ? s_idx[lane] : 0xffffffffu;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            float ov = __shfl_down_sync(0xffffffffu, best, off);
            uint32_t oi = __shfl_down_sync(0xffffffffu, bidx, off);
            if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
        }
        if (lane == 0) { PeakRec pr; pr.idx = bidx; pr.val = best; peaks[(size_t)b * nch + c] = pr; }
    }
}

cudaError_t launch_correlate_p2(const CorrLaunch &p, cudaStream_t st) {
    static bool attr_done[64] = {};
    int dev = 0; cudaGetDevice(&dev);
    if (!attr_done[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(k_correlate_p2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P13::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done[dev & 63] = true;
    }
    if (p.n_items)
        k_correlate_p2<<<p.n_items, kT2, P13::kSmemBytes, st>>>(
            p.iq, p.nch, p.B, p.item_list, p.nch - 1, p.block_slot, p.xormask, p.TN,
            reinterpret_cast<const float4 *>(p.refspec), p.peaks);
    return cudaGetLastError();
}

}  // namespace coh
