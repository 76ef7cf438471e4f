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
#include "tma.cuh"

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

// ---- last inverse pass + arg-max of one thread ------------------------------------------------------
// last inverse pass (radix 16, stride L/16) kept in registers; recombine the halves,
// |.|^2 (cdsp::magsquared, src/cdsp.cc:100-103) and running arg-max with lowest-index ties
// (volk_32f_index_max_32u semantics, src/cdsp.cc:135-139)
template <int LOG2L, int PADSH = 4>
__device__ __forceinline__ void last_pass_argmax(const float4 *smem4, float2 w_l, float2 w_n, float &best, uint32_t &bidx) {
    using P = Plan<LOG2L>;
    constexpr int L = P::L, S = P::S1;
    for (int low = threadIdx.x; low < S; low += kFftThreads) {
        PC v[16];
        {
            const TwPair tw(cconj(w_l), cconj(w_n));
#pragma unroll
            for (int k = 0; k < 16; k++) v[k] = tw.apply(pc_from(smem4[ppad_t<PADSH>(low + k * S)]), k);
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
}

// (val, idx) arg-max across the lanes of a warp, lowest index on ties
__device__ __forceinline__ void warp_argmax(float &best, uint32_t &bidx) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        float ov = __shfl_down_sync(0xffffffffu, best, off);
        uint32_t oi = __shfl_down_sync(0xffffffffu, bidx, off);
        if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
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

    float best = -1.0f; uint32_t bidx = 0xffffffffu;
    last_pass_argmax<LOG2L>(smem4, w_l, w_n, best, bidx);
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

// ---- K2, persistent form for L = 8192 ----------------------------------------------------------------
// One CTA per SM draws (block, channel) items through a device ticket.  Per item the shared-memory
// transform has two cross-warp transposes (pass 1 writes element `low` of all 16 sub-transforms, the last
// inverse pass reads it back) and five warp-local stages in between.  A thread's last-pass reads and its
// pass-1 writes touch the SAME sixteen slots {low + k L/16}, so the last pass of item i-1 and pass 1 of
// item i run back to back in one phase (X) with no barrier between them:
//     X(i): last pass + arg-max of item i-1, then pass 1 of item i      -- barrier
//     M(i): forward mid pass, fused 32-point stage with x conj(R), inverse mid pass (warp local) -- barrier
// The CTA never drains between items (no launch gap, no cold start on pass 1's loads, two block barriers
// per item instead of three), and the raw bytes of item i+1 arrive by ONE bulk asynchronous copy (TMA,
// cp.async.bulk + mbarrier) into a 16 KiB staging buffer while item i is transformed -- pass 1 reads
// them with LDS instead of sixteen 2-byte global loads per thread.
struct ItemMeta { uint32_t b, c, slot, valid; };
constexpr size_t kPersistSmemData = ((size_t)ppadded_len_t<5>(8192) * sizeof(float4) + 127) / 128 * 128;

// reference spectrum in the register order of the fused 32-point stage (fft_passes.cuh, fused_b_out)
template <int LOG2L>
__global__ void __launch_bounds__(kFftThreads, 1)
k_ref_spectrum_fused(const uint8_t *__restrict__ iq, size_t block_stride, const uint32_t *__restrict__ block_list,
                     uint32_t xormask, const float2 *__restrict__ TN, float4 *__restrict__ refspec) {
    using P = Plan<LOG2L>;
    extern __shared__ __align__(16) float4 smem4[];
    const uint32_t slot = blockIdx.x;
    const uint32_t b = block_list ? block_list[slot] : slot;
    const uint8_t *x = iq + (size_t)b * block_stride;          // channel 0 of block b
    const float2 w_a = mid_twiddle<LOG2L, P::RA, P::SA>(TN);
    pass1_butterfly<LOG2L, true, 5>(x, xormask, smem4, threadIdx.x, __ldg(&TN[2 * threadIdx.x]), __ldg(&TN[threadIdx.x]));
    __syncthreads();
    mid_pass<LOG2L, P::RA, P::SA, -1, 5>(smem4, w_a);
    __syncwarp();
    fused_b_out<LOG2L>(smem4, refspec + (size_t)slot * P::L);
}

template <int LOG2L>
__global__ void __launch_bounds__(kFftThreads, 1)
k_correlate_persist(const uint8_t *__restrict__ iq, uint32_t nch, uint32_t B, const uint32_t *__restrict__ item_list,
                    uint32_t n_signal, uint32_t n_items, const uint32_t *__restrict__ block_slot, uint32_t xormask,
                    const float2 *__restrict__ TN, const float4 *__restrict__ refspec, PeakRec *__restrict__ peaks,
                    uint32_t *__restrict__ ticket) {
    using P = Plan<LOG2L>;
    constexpr int L = P::L, S = P::S1;
    static_assert(S == kFftThreads, "one pass-1 / last-pass butterfly per thread");
    extern __shared__ __align__(16) float4 smem4[];
    uint8_t *raw0 = reinterpret_cast<uint8_t *>(smem4) + kPersistSmemData;   // 2 x B staging
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ ItemMeta s_meta[2];
    __shared__ float s_val[kFftThreads / 32];
    __shared__ uint32_t s_idx[kFftThreads / 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    float2 w_l = __ldg(&TN[2 * threadIdx.x]), w_n = __ldg(&TN[threadIdx.x]);   // W_L^low, W_N^low
    float2 w_a = mid_twiddle<LOG2L, P::RA, P::SA>(TN);                          // W_512^(lane): mid pass A
    float2 t32 = make_float2(kW32c[lane & 15], kW32s[lane & 15]);               // conj(W_32^(lane & 15)): inverse 32-point stage

    // thread 0 owns the ticket: `tk` is the ticket of the item AFTER the one being staged, drawn one
    // phase ahead so that its latency is never waited for
    uint32_t tk = 0;
    auto stage = [&](uint32_t t, int buf) {                 // thread 0: describe item t, start its copy
        ItemMeta m; m.valid = t < n_items ? 1u : 0u; m.b = m.c = m.slot = 0;
        if (m.valid) {
            if (item_list) { const uint32_t it = item_list[t]; m.b = it / nch; m.c = it % nch; }
            else { m.b = t / n_signal; m.c = 1 + t % n_signal; }
            m.slot = block_slot ? block_slot[m.b] : m.b;
            tma_row_load(smem_u32(raw0 + (size_t)buf * B), iq + ((size_t)m.b * nch + m.c) * B, B, smem_u32(&s_bar[buf]));
        }
        s_meta[buf] = m;
    };
    if (threadIdx.x == 0) {
        mbar_init(smem_u32(&s_bar[0]), 1); mbar_init(smem_u32(&s_bar[1]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        stage(atomicAdd(ticket, 1u), 0);
        // a second item is drawn ahead only when there is more than one wave of them (a single block of a
        // live stream has fewer items than SMs: every CTA takes exactly one)
        tk = n_items > gridDim.x ? atomicAdd(ticket, 1u) : n_items;
    }
    __syncthreads();

    ItemMeta prev; prev.valid = 0; prev.b = prev.c = prev.slot = 0;
    for (uint32_t i = 0;; i++) {
        const int buf = (int)(i & 1u);
        const ItemMeta cur = s_meta[buf];
        // the twiddle bases are loop invariant, and so is every power the passes derive from them: left alone,
        // ptxas hoists ~220 words of derived twiddles per thread out of the loop and spills them to local
        // memory.  Recomputing them per item (as the one-item-per-CTA kernel does) is cheaper than that traffic.
#ifndef COH_CORR_HOIST
        asm volatile("" : "+f"(w_l.x), "+f"(w_l.y), "+f"(w_n.x), "+f"(w_n.y));
        asm volatile("" : "+f"(w_a.x), "+f"(w_a.y), "+f"(t32.x), "+f"(t32.y));
#endif
        if (threadIdx.x == 0 && cur.valid) {                // item i+1: its staging buffer was last read in X(i-1)
            stage(tk, buf ^ 1);
            if (tk < n_items) tk = atomicAdd(ticket, 1u);
        }
        // ---- X(i) -------------------------------------------------------------------------------
        if (prev.valid) {
            float best = -1.0f; uint32_t bidx = 0xffffffffu;
            last_pass_argmax<LOG2L, 5>(smem4, w_l, w_n, best, bidx);
            warp_argmax(best, bidx);
            if (lane == 0) { s_val[warp] = best; s_idx[warp] = bidx; }
        }
#ifdef COH_EARLY_BEFORE_BAR
        float4 r_early[kEarly ? kEarly : 1];
#endif
        if (cur.valid) {
            mbar_wait(smem_u32(&s_bar[buf]), (i >> 1) & 1u);
            pass1_butterfly<LOG2L, false, 5>(raw0 + (size_t)buf * B, xormask, smem4, threadIdx.x, w_l, w_n);
#ifdef COH_EARLY_BEFORE_BAR
            // the first spectrum rows do not depend on shared memory: requested before the block barrier, their L2
            // latency would overlap the wait for the slowest warp.  MEASURED SLOWER (round 3, B200): 1.719 -> 1.897 ms
            // per 21 504 items (1.936 with COH_R_EARLY=4) -- eight float4 live across the barrier push the kernel
            // deeper into its 128-register cap (spill loads 56 -> 96 B) and the rows of the fast warps arrive
            // long before they are used.  Kept as a build variant only (profiles/r3_correlate_history.md).
            fused_b_early<LOG2L>(refspec + (size_t)cur.slot * L, r_early);
#endif
        }
        __syncthreads();
        if (prev.valid && warp == 0) {
            float best = lane < kFftThreads / 32 ? s_val[lane] : -1.0f;
            uint32_t bidx = lane < kFftThreads / 32 ? s_idx[lane] : 0xffffffffu;
            warp_argmax(best, bidx);
            if (lane == 0) { PeakRec pr; pr.idx = bidx; pr.val = best; peaks[(size_t)prev.b * nch + prev.c] = pr; }
        }
        if (!cur.valid) break;
        // ---- M(i) -------------------------------------------------------------------------------
#ifndef COH_EARLY_BEFORE_BAR
        float4 r_early[kEarly ? kEarly : 1];
        fused_b_early<LOG2L>(refspec + (size_t)cur.slot * L, r_early);
#endif
        mid_pass<LOG2L, P::RA, P::SA, -1, 5>(smem4, w_a);          // radix 16 at stride 32
        __syncwarp();
        fused_b_stage<LOG2L>(smem4, refspec + (size_t)cur.slot * L, r_early);   // 32-point stage, x conj(R), inverse 16-point DFT
        __syncwarp();
        inv_a_stage<LOG2L>(smem4, w_a, t32);                       // radix 16 at stride 32, decimation in time
        __syncthreads();
        prev = cur;
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
    bool fused = false;
#ifndef COH_CORRELATE_V1
    fused = LOG2L == 13 && p.ticket != nullptr;          // persistent kernel + its reference-spectrum layout
#endif
    if (p.n_ref_blocks) {
        if constexpr (LOG2L == 13) {
            if (fused) {
                static bool attr3[64] = {};
                if (!attr3[dev & 63]) {
                    cudaError_t e = cudaFuncSetAttribute(k_ref_spectrum_fused<LOG2L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPersistSmemData);
                    if (e != cudaSuccess) return e;
                    attr3[dev & 63] = true;
                }
                k_ref_spectrum_fused<LOG2L><<<p.n_ref_blocks, kFftThreads, kPersistSmemData, st>>>(
                    p.iq, (size_t)p.nch * p.B, p.ref_block_list, p.xormask, p.TN, reinterpret_cast<float4 *>(p.refspec));
            }
        }
        if (!fused)
            k_ref_spectrum<LOG2L><<<p.n_ref_blocks, kFftThreads, P::kSmemBytes, st>>>(
                p.iq, (size_t)p.nch * p.B, p.ref_block_list, p.xormask, p.TN, reinterpret_cast<float4 *>(p.refspec));
    }
    if (p.n_items) {
#ifndef COH_CORRELATE_V1
        if constexpr (LOG2L == 13) {
            if (p.ticket) {
                const size_t smem = kPersistSmemData + 2 * (size_t)p.B;
                static bool attr2[64] = {};
                if (!attr2[dev & 63]) {
                    cudaError_t e = cudaFuncSetAttribute(k_correlate_persist<LOG2L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                    if (e != cudaSuccess) return e;
                    attr2[dev & 63] = true;
                }
                cudaError_t e = cudaMemsetAsync(p.ticket, 0, sizeof(uint32_t), st);
                if (e != cudaSuccess) return e;
                const uint32_t sms = (uint32_t)(p.sm_count > 0 ? p.sm_count : 148);
                const uint32_t grid = p.n_items < sms ? p.n_items : sms;
                k_correlate_persist<LOG2L><<<grid, kFftThreads, smem, st>>>(
                    p.iq, p.nch, p.B, p.item_list, p.nch - 1, p.n_items, p.block_slot, p.xormask, p.TN,
                    reinterpret_cast<const float4 *>(p.refspec), p.peaks, p.ticket);
                return cudaGetLastError();
            }
        }
#endif
        k_correlate<LOG2L><<<p.n_items, kFftThreads, P::kSmemBytes, st>>>(
            p.iq, p.nch, p.B, p.item_list, p.nch - 1, p.block_slot, p.xormask, p.TN, reinterpret_cast<const float4 *>(p.refspec), p.peaks);
    }
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
