// fft_passes.cuh -- the shared-memory passes of the packed two-half transform, used by the
// single-CTA correlation kernels (fft_kernels.cu, L <= 8192) and by the split large-block path
// (fft_big.cu, L up to 65536).  See fft_kernels.cu for the transform layout.
#pragma once
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
template <int LOG2L, bool NEG_B, int PADSH = 4>
__device__ __forceinline__ void pass1_butterfly(const uint8_t *x, uint32_t xormask, float4 *s, int low,
                                                float2 w_l, float2 w_n) {
    constexpr int S = Plan<LOG2L>::S1;
    PC v[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const float2 a = load_sample(x, low + k * S, xormask);
        const float2 b = mul_w32<-1>(a, k);           // W_N^(k*L/16) = W_32^k
        v[k].r = make_float2(a.x, b.x); v[k].i = make_float2(a.y, b.y);
    }
    PDft<16, -1>::run(v);
    float2 g = w_n;                                   // W_N^low (odd-bin modulation)
    if (NEG_B) g = make_float2(-g.x, -g.y);           // ref sits in [L,2L): odd bins change sign
    const TwPair tw(w_l, g);                          // W_L^low
#pragma unroll
    for (int k = 0; k < 16; k++) s[ppad_t<PADSH>(low + k * S)] = pc_to(tw.apply(v[k], k));
}
template <int LOG2L, bool NEG_B>
__device__ __forceinline__ void pass1_from_global(const uint8_t *x, uint32_t xormask, float4 *s,
                                                  const float2 *__restrict__ TN) {
    constexpr int S = Plan<LOG2L>::S1;
    for (int low = threadIdx.x; low < S; low += kFftThreads)
        pass1_butterfly<LOG2L, NEG_B>(x, xormask, s, low, __ldg(&TN[2 * low]), __ldg(&TN[low]));
}
// one butterfly per thread, twiddle bases already in registers
template <int LOG2L, bool NEG_B>
__device__ __forceinline__ void pass1_from_global(const uint8_t *x, uint32_t xormask, float4 *s,
                                                  float2 w_l, float2 w_n) {
    if ((int)threadIdx.x < Plan<LOG2L>::S1) pass1_butterfly<LOG2L, NEG_B>(x, xormask, s, threadIdx.x, w_l, w_n);
}

// ---- in-place mid passes ----------------------------------------------------------------------
// Every mid pass has at most one butterfly per thread (L <= 8192, 512 threads), so its twiddle base
// W_(S*R)^low depends on the thread only: mid_twiddle() fetches it once, ahead of the barriers,
// instead of a dependent global load at the top of each pass.
template <int LOG2L, int R, int S>
__device__ __forceinline__ float2 mid_twiddle(const float2 *__restrict__ TN) {
    using P = Plan<LOG2L>;
    constexpr int NB = P::L / R;                      // butterflies
    constexpr int TSTEP = 2 * (P::L / (S * R));       // index step into TN for W_(S*R)
    static_assert(NB <= kFftThreads, "one butterfly per thread");
    const int bf = threadIdx.x < NB ? threadIdx.x : 0;
    return __ldg(&TN[(bf & (S - 1)) * TSTEP]);
}
struct MidTw { float2 a, b; };
template <int LOG2L>
__device__ __forceinline__ MidTw mid_twiddles(const float2 *__restrict__ TN) {
    using P = Plan<LOG2L>;
    MidTw t; t.a = t.b = make_float2(1.0f, 0.0f);
    if constexpr (P::RA > 1) t.a = mid_twiddle<LOG2L, P::RA, P::SA>(TN);
    if constexpr (P::RB > 1) t.b = mid_twiddle<LOG2L, P::RB, 2>(TN);
    return t;
}

template <int LOG2L, int R, int S, int DIR, int PADSH = 4>
__device__ __forceinline__ void mid_pass(float4 *s, float2 w) {
    using P = Plan<LOG2L>;
    constexpr int NB = P::L / R;                      // butterflies
    if (DIR > 0) w = cconj(w);
    {
        const int bf = threadIdx.x;
        if (bf >= NB) return;
        const int low = bf & (S - 1), high = bf / S;
        const int base = high * S * R + low;
        const TwScalar tw(w);
        PC v[R];
#pragma unroll
        for (int k = 0; k < R; k++) v[k] = pc_from(s[ppad_t<PADSH>(base + k * S)]);
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
        for (int k = 0; k < R; k++) s[ppad_t<PADSH>(base + k * S)] = pc_to(v[k]);
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
__device__ __forceinline__ void mid_passes(float4 *s, const MidTw tw) {
    using P = Plan<LOG2L>;
    if (DIR < 0) {
        if constexpr (P::RA > 1) { mid_pass<LOG2L, P::RA, P::SA, -1>(s, tw.a); mid_sync<LOG2L>(); }
        if constexpr (P::RB > 1) { mid_pass<LOG2L, P::RB, 2, -1>(s, tw.b); mid_sync<LOG2L>(); }
    } else {
        if constexpr (P::RB > 1) { mid_pass<LOG2L, P::RB, 2, +1>(s, tw.b); mid_sync<LOG2L>(); }
        if constexpr (P::RA > 1) { mid_pass<LOG2L, P::RA, P::SA, +1>(s, tw.a); mid_sync<LOG2L>(); }
    }
}
template <int LOG2L, int DIR>
__device__ __forceinline__ void mid_passes(float4 *s, const float2 *TN) {
    mid_passes<LOG2L, DIR>(s, mid_twiddles<LOG2L>(TN));
}

// ---- pair stage --------------------------------------------------------------------------------
// last forward radix-2 stage written out: the spectrum in the digit-reversed order of the passes
template <int LOG2L>
__device__ __forceinline__ void pair_out(const float4 *s, float4 *__restrict__ out) {
    using P = Plan<LOG2L>;
    constexpr int kTrips = (P::L / 2 + kFftThreads - 1) / kFftThreads;
    for (int j = 0; j < kTrips; j++) {
        const int m = pair_index<LOG2L>(j);
        if (m >= P::L / 2) break;
        const PC e0 = pc_from(s[ppad(2 * m)]), e1 = pc_from(s[ppad(2 * m + 1)]);
        out[2 * m] = pc_to(pc_add(e0, e1));
        out[2 * m + 1] = pc_to(pc_sub(e0, e1));
    }
}

// last forward radix-2 stage, X * conj(R) (cdsp::conjugatemul, src/cdsp.cc:105-108), first inverse
// radix-2 stage -- all on adjacent pairs, in registers
template <int LOG2L>
__device__ __forceinline__ void pair_correlate(float4 *s, const float4 *__restrict__ R) {
    using P = Plan<LOG2L>;
    constexpr int kTrips = (P::L / 2 + kFftThreads - 1) / kFftThreads;
    constexpr int kGroup = kTrips >= 4 ? 4 : kTrips;      // spectrum loads of a whole group in flight
    // lanes 0-3 of every quarter-warp touch the even element first, lanes 4-7 the odd one:
    // the eight 16-byte words of one LDS.128 / STS.128 phase then fall on distinct bank groups
    const int sw = (threadIdx.x >> 2) & 1;
#pragma unroll 1
    for (int j0 = 0; j0 < kTrips; j0 += kGroup) {
        float4 g0[kGroup], g1[kGroup];
#pragma unroll
        for (int u = 0; u < kGroup; u++) {
            const int m = pair_index<LOG2L>(j0 + u);
            if (j0 + u < kTrips && m < P::L / 2) { g0[u] = __ldg(&R[2 * m]); g1[u] = __ldg(&R[2 * m + 1]); }
        }
#pragma unroll
        for (int u = 0; u < kGroup; u++) {
            const int m = pair_index<LOG2L>(j0 + u);
            if (j0 + u >= kTrips || m >= P::L / 2) continue;
            const float4 q0 = s[ppad(2 * m + sw)], q1 = s[ppad(2 * m + 1 - sw)];
            const PC e0 = pc_from(sw ? q1 : q0), e1 = pc_from(sw ? q0 : q1);
            const PC r0 = pc_from(g0[u]), r1 = pc_from(g1[u]);
            const PC z0 = pc_mulcp(pc_add(e0, e1), r0), z1 = pc_mulcp(pc_sub(e0, e1), r1);
            const float4 o0 = pc_to(pc_add(z0, z1)), o1 = pc_to(pc_sub(z0, z1));
            s[ppad(2 * m + sw)] = sw ? o1 : o0;
            s[ppad(2 * m + 1 - sw)] = sw ? o0 : o1;
        }
    }
}

// ---- fused 32-point stage of the persistent L = 8192 kernel -------------------------------------------
// After pass 1 (radix 16, stride 512) and mid pass A (radix 16, stride 32) the data is 256 contiguous
// 32-point sub-transforms, 16 per warp.  Instead of radix 16 at stride 2 + a pair stage (two more
// shared-memory round trips, the second with hardly any arithmetic between its loads and stores -- 20 % of
// the kernel's time in profiles/r2_ncu_k_correlate_persist_v1.txt), TWO lanes share one 32-point group g:
// both read all 32 elements (the lanes l and l + 16 of a warp read the same addresses), lane `half` = 0
// forms the sums s[j] = e[j] + e[j+16], lane `half` = 1 the differences d[j] = (e[j] - e[j+16]) W_32^j, and
// each finishes with a 16-point DFT in registers: the even / odd bins of the group.  The product with the
// conjugate reference spectrum (cdsp::conjugatemul, src/cdsp.cc:105-108) and the first inverse 16-point DFT
// follow in the same registers; what is stored back are the two half-results E and O, which the inverse mid
// pass A combines on load (inv_a_stage): y[j] = E[j] + conj(W_32^j) O[j], y[j+16] = E[j] - conj(W_32^j) O[j].
// Shared-memory padding for these passes: one element per 32 (ppad_t<5>), which keeps the stride-33 group
// accesses of a quarter-warp on distinct 16-byte bank groups.
//
// Reference-spectrum layout for this path: bin n (0..15) of lane `lane` of warp w at
// R[(w * 16 + n) * 32 + lane] -- k_ref_spectrum runs the same forward code and stores its registers, so the
// order is consistent by construction and every load is one coalesced 512-byte row.
template <int LOG2L>
__device__ __forceinline__ void fused_b_load(const float4 *s, PC (&v)[16]) {
    static_assert(LOG2L == 13 && kFftThreads == 512, "16 groups of 32 points per warp");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int half = lane >> 4, base = warp * 512 + (lane & 15) * 32;
    const float2 sg = splat(half ? -1.0f : 1.0f);
#pragma unroll
    for (int j = 0; j < 16; j++) {
        const PC lo = pc_from(s[ppad_t<5>(base + j)]), hi = pc_from(s[ppad_t<5>(base + 16 + j)]);
        PC d; d.r = __ffma2_rn(hi.r, sg, lo.r); d.i = __ffma2_rn(hi.i, sg, lo.i);   // e[j] +- e[j+16]
        if (j == 0) { v[j] = d; continue; }
        // W_32^j = exp(-2 pi i j / 32) on the difference lane, 1 on the sum lane
        const float c = half ? kW32c[j] : 1.0f, sn = half ? -kW32s[j] : 0.0f;
        v[j] = pc_mul(d, make_float2(c, sn));
    }
    PDft<16, -1>::run(v);
}

// k_ref_spectrum: the registers of fused_b_load are the spectrum
template <int LOG2L>
__device__ __forceinline__ void fused_b_out(const float4 *s, float4 *__restrict__ R) {
    PC v[16];
    fused_b_load<LOG2L>(s, v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int n = 0; n < 16; n++) R[(warp * 16 + n) * 32 + lane] = pc_to(v[n]);
}

// k_correlate: forward 32-point stage, x conj(R), first inverse 16-point DFT.  The spectrum rows come from L2
// (one 128 KiB spectrum per block, shared by the block's channels): kEarly of this lane's sixteen rows are
// requested by the caller ahead of mid pass A (r_early), the next 8 - kEarly before the 32-point stage's own
// shared-memory loads, the last eight behind the DFT.
#ifndef COH_R_EARLY
#define COH_R_EARLY 8       // measured 0 / 4 / 8: 1.735 / 1.772 / 1.717 ms per 21 504 items (profiles/r2_correlate_history.md)
#endif
constexpr int kEarly = COH_R_EARLY;
template <int LOG2L>
__device__ __forceinline__ void fused_b_early(const float4 *__restrict__ R, float4 (&re)[kEarly ? kEarly : 1]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float4 *Rw = R + (warp * 16) * 32 + lane;
#pragma unroll
    for (int n = 0; n < kEarly; n++) re[n] = __ldg(Rw + n * 32);
}
template <int LOG2L>
__device__ __forceinline__ void fused_b_stage(float4 *s, const float4 *__restrict__ R, const float4 (&re)[kEarly ? kEarly : 1]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float4 *Rw = R + (warp * 16) * 32 + lane;
    float4 r0[8];                                       // first half of the spectrum rows in flight across the DFT
#pragma unroll
    for (int n = 0; n < 8; n++) r0[n] = n < kEarly ? re[n] : __ldg(Rw + n * 32);
    PC v[16];
    fused_b_load<LOG2L>(s, v);
    float4 r1[8];
#pragma unroll
    for (int n = 0; n < 8; n++) r1[n] = __ldg(Rw + (8 + n) * 32);
#pragma unroll
    for (int n = 0; n < 8; n++) v[n] = pc_mulcp(v[n], pc_from(r0[n]));
#pragma unroll
    for (int n = 0; n < 8; n++) v[8 + n] = pc_mulcp(v[8 + n], pc_from(r1[n]));
    PDft<16, +1>::run(v);
    const int base = warp * 512 + (lane & 15) * 32 + (lane >> 4) * 16;
    __syncwarp();                                       // the partner lane has read what is overwritten now
#pragma unroll
    for (int j = 0; j < 16; j++) s[ppad_t<5>(base + j)] = pc_to(v[j]);
}

// inverse mid pass A (radix 16, stride 32, decimation in time) reading the E / O halves the fused stage left
template <int LOG2L>
__device__ __forceinline__ void inv_a_stage(float4 *s, float2 w, float2 t32) {
    // w = W_512^low (forward twiddle base of mid pass A), t32 = conj(W_32^(low & 15))
    const int warp = threadIdx.x >> 5, low = threadIdx.x & 31;
    const int sub = warp * 512, jl = low & 15;
    const float2 sg = splat((low >> 4) ? -1.0f : 1.0f);
    const TwScalar tw(cconj(w));
    PC v[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const PC E = pc_from(s[ppad_t<5>(sub + 32 * k + jl)]), O = pc_from(s[ppad_t<5>(sub + 32 * k + 16 + jl)]);
        const PC t = pc_mul(O, t32);
        PC y; y.r = __ffma2_rn(t.r, sg, E.r); y.i = __ffma2_rn(t.i, sg, E.i);
        v[k] = k ? tw.apply(y, k) : y;
    }
    PDft<16, +1>::run(v);
    __syncwarp();                                       // lanes low and low ^ 16 read each other's slots
#pragma unroll
    for (int k = 0; k < 16; k++) s[ppad_t<5>(sub + low + 32 * k)] = pc_to(v[k]);
}

}  // namespace coh
