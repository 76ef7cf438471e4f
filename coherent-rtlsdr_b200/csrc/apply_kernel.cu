// apply_kernel.cu -- K3: shift, phase estimate, phase IIR, rotation and int8 packing, written
// straight into the packetizer's column layout (sm_100a).  In TRACK mode (no lag requested,
// the reference's steady state) this is the only kernel of the path and it is HBM-bound:
// 2 bytes read + 2 bytes written per complex sample.
//
// Replaces, in the reference tree, per signal channel and block (src/ccoherent.cc:262-283):
//   cdsp::convtosigned          src/cdsp.cc:21-34 (via cbuffer::setbufferptr, include/common.h:121)
//   crtlsdr::convtofloat        src/crtlsdr.cc:205-207 -> volk_8i_s32f_convert_32f
//   csdrdevice::est_phasecorrect src/csdrdevice.cc:58-69 (conjugate dot product + IIR)
//   csdrdevice::phasecorrect    src/csdrdevice.cc:80-84 -> cdsp::scalarmul src/cdsp.cc:46-49
//   cpacketize::write           src/cpacketizer.cc:137-172 -> cdsp::convto8bit src/cdsp.cc:51-54
// plus the software integer shift the reference leaves to the dongles' resamplers.
//
// One WARP handles one (block, channel) item end to end and shares nothing with the other warps
// of its CTA (no block barrier): lane 0 fetches the channel row -- 16 KiB segment by segment -- into
// the warp's private shared-memory buffer with ONE bulk asynchronous copy (TMA, cp.async.bulk +
// mbarrier), the warp forms the dot product from shared memory, looks back, and rotates / packs
// the same resident bytes (rows longer than a segment are re-fetched, from L2).  Thirteen such
// warps per SM overlap each other's fetch and look-back latencies, which a CTA-per-item design
// with block barriers could not (profiles/r1_apply_history.md).  Warps are persistent and draw
// items through an atomic ticket in (block, channel) order, only when they are ready to start one;
// every item with a smaller ticket is therefore already being processed by a resident warp that
// never waits on a larger one -- which makes the look-back waits below deadlock-free.  The warp
// that draws the last ticket of a launch resets the counter.
//
// Phase estimate: the samples are int8-valued, so S = sum x'[n] * conj(r[n]) is accumulated
// exactly in integers with dp4a; only the direction conj(S)/|S| is taken in float.  The IIR
// p_b = alpha*q_b + (1-alpha)*p_(b-1) couples consecutive blocks of a channel; each item
// publishes its raw estimate q_b and evaluates the recurrence over the previous `lookback`
// published estimates (decoupled look-back; the truncated tail weighs (1-alpha)^lookback).
// A published record is ONE aligned 16-byte word {q.re, q.im, epoch, check}: value and ready
// flag travel in a single transaction, so no fence (and no L1 invalidation) is needed.
//
// Arithmetic runs two samples at a time on Blackwell's packed fp32 pipe (FADD2/FMUL2/FFMA2)
// with the real parts of a word in one register pair and the imaginary parts in the other.
#include "kernels.h"
#include "tma.cuh"

namespace coh {

#ifndef COH_APPLY_WARPS
#define COH_APPLY_WARPS 13
#endif
#ifndef COH_APPLY_SEG
#define COH_APPLY_SEG 16384
#endif
constexpr int kWarpsPerCta = COH_APPLY_WARPS;             // independent warps, one CTA per SM
constexpr int kApplyThreads = 32 * kWarpsPerCta;
constexpr uint32_t kSegBytes = COH_APPLY_SEG;             // shared-memory resident part of a row
constexpr uint32_t kScratchBytes = 512;                   // look-back scratch: 64 x float2
// + one chunk of overlap for the shift; padded to a multiple of 128 B so that every warp's buffer starts on a
// bank-row boundary (a pitch of 16928 = 32 mod 128 made 12 of 13 warps' LDS.128 bursts straddle five bank rows:
// 2.67 M load conflicts per step in profiles/r1e_ncu_k_apply.txt)
#ifndef COH_APPLY_PAD
#define COH_APPLY_PAD 128
#endif
constexpr uint32_t kBufBytes = (kSegBytes + 32 + kScratchBytes + COH_APPLY_PAD - 1) / COH_APPLY_PAD * COH_APPLY_PAD;
constexpr int kU = 2;                                     // 16-byte chunks per lane per loop trip
constexpr int kMaxLookback = 2048;

// ---- 16-byte publish / poll ------------------------------------------------------------------
__device__ __forceinline__ void publish16(uint4 *p, uint32_t a, uint32_t b, uint32_t epoch) {
    uint32_t chk = a ^ b ^ epoch ^ 0x5bd1e995u;
    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1, %2, %3, %4};"
                 ::"l"(p), "r"(a), "r"(b), "r"(epoch), "r"(chk) : "memory");
}
__device__ __forceinline__ uint4 poll16(const uint4 *p, uint32_t epoch) {
    uint4 v;
    for (;;) {
        asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
        if (v.z == epoch && v.w == (v.x ^ v.y ^ epoch ^ 0x5bd1e995u)) return v;
        __nanosleep(40);
    }
}

// bulk asynchronous copy (TMA) + mbarrier, one barrier per warp: tma.cuh

// ---- shifted read: x'[n] = x[n + sh] (input domain, zero_code outside the block) --------------
// The byte offset 16*j + 2*sh has the same residue mod 16 for every 16-byte chunk j, so the
// realignment (word select + 16-bit funnel shift) is uniform across the item: the dot and the
// rotate loops exist in eight straight-line variants, selected once per item.
template <int K>   // K = (2*sh & 15) / 2 in 0..7: out = bytes [2K, 2K+16) of the pair (A, B)
__device__ __forceinline__ uint4 realign(const uint4 &A, const uint4 &Bv) {
    if (K == 0) return A;
    const uint32_t W[8] = { A.x, A.y, A.z, A.w, Bv.x, Bv.y, Bv.z, Bv.w };
    constexpr int ws = K >> 1;
    if (K & 1)
        return make_uint4(__funnelshift_r(W[ws], W[ws + 1], 16), __funnelshift_r(W[ws + 1], W[ws + 2], 16),
                          __funnelshift_r(W[ws + 2], W[ws + 3], 16), __funnelshift_r(W[ws + 3], W[ws + 4], 16));
    return make_uint4(W[ws], W[ws + 1], W[ws + 2], W[ws + 3]);
}

// chunk jj of the segment (row chunk j = seg0 + jj) out of the resident buffer; source chunks
// a = j + a0 and a + 1 sit at buffer chunks jj and jj + 1; chunks outside the row read as ZERO
template <int K, uint32_t ZERO>
__device__ __forceinline__ uint4 read_shifted(const uint4 *sx, int jj, int a, int nchunks) {
    const uint4 z = make_uint4(ZERO, ZERO, ZERO, ZERO);
    const uint4 A = (unsigned)a < (unsigned)nchunks ? sx[jj] : z;
    if (K == 0) return A;
    const uint4 Bv = (unsigned)(a + 1) < (unsigned)nchunks ? sx[jj + 1] : z;
    return realign<K>(A, Bv);
}

// ---- one 32-bit word = two complex samples ----------------------------------------------------
struct RotConst { float2 px, py, npy; };    // (p.re,p.re), (p.im,p.im), (-p.im,-p.im)

// xu: offset-binary bytes [r0 i0 r1 i1].  Returns packed signed bytes of sat8(rint(127 * x*p)).
//   convert : volk_8i_s32f_convert_32f   -> (float)s8 * float(1/127)            (cdsp.cc:41-44)
//   rotate  : volk_32fc_s32fc_multiply_32fc, AVX+FMA kernel rounding
//             re = fma(xr, p.re, -(xi*p.im)), im = fma(xi, p.re, xr*p.im)        (cdsp.cc:46-49)
//   pack    : volk_32f_s32f_convert_8i   -> sat8(rint(v * 127))                  (cdsp.cc:51-54)
template <bool CF>
__device__ __forceinline__ uint32_t rotate_pack_word(uint32_t xu, const RotConst &k, float2 *cf_out) {
    float2 R = make_float2(__uint_as_float(__byte_perm(xu, 0x4B000000u, 0x7440)),
                           __uint_as_float(__byte_perm(xu, 0x4B000000u, 0x7442)));
    float2 I = make_float2(__uint_as_float(__byte_perm(xu, 0x4B000000u, 0x7441)),
                           __uint_as_float(__byte_perm(xu, 0x4B000000u, 0x7443)));
    const float2 mk = make_float2(-8388736.0f, -8388736.0f);            // -(2^23 + 128)
    const float2 c127 = make_float2(0x1.020408p-7f, 0x1.020408p-7f);   // float(1.0/127.0)
    const float2 s127 = make_float2(127.0f, 127.0f);
    R = __fmul2_rn(__fadd2_rn(R, mk), c127);
    I = __fmul2_rn(__fadd2_rn(I, mk), c127);
    float2 yR = __ffma2_rn(R, k.px, __fmul2_rn(I, k.npy));
    float2 yI = __ffma2_rn(I, k.px, __fmul2_rn(R, k.py));
    if (CF) { cf_out[0] = make_float2(yR.x, yI.x); cf_out[1] = make_float2(yR.y, yI.y); }
    float2 sR = __fmul2_rn(yR, s127), sI = __fmul2_rn(yI, s127);
    int r0 = __float2int_rn(sR.x), r1 = __float2int_rn(sR.y);           // NaN -> 0
    int i0 = __float2int_rn(sI.x), i1 = __float2int_rn(sI.y);
    uint32_t hi, out;
    asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(hi) : "r"(i1), "r"(r1), "r"(0));
    asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(out) : "r"(i0), "r"(r0), "r"(hi));
    return out;
}


// ---- pass 1 over one resident segment: exact integer conjugate dot product ----------------------
// S = sum x'[n] conj(r[n]) on the signed bytes; acc = {sum xr*rr + xi*ri, sum xi*rr + xr*ri, sum xr*ri}.
// The shifted, zero-filled chunks x' are written back over the buffer (in place: a trip reads
// chunks [jj0, jj0 + 64] and, after a warp sync, writes [jj0, jj0 + 64), which no later trip reads
// as its overlap chunk), so that pass 2 exists in one variant and reads aligned 16-byte words.
template <int K, uint32_t TOSIGNED>
__device__ __forceinline__ void dot_segment(uint4 *sx, const uint4 *__restrict__ rin, int seg_chunks,
                                            int a_base, int nchunks, int lane, int (&acc)[3]) {
    const uint4 z = make_uint4(TOSIGNED, TOSIGNED, TOSIGNED, TOSIGNED);
    uint4 rv[kU];
#pragma unroll
    for (int u = 0; u < kU; u++) rv[u] = lane + 32 * u < seg_chunks ? __ldg(rin + lane + 32 * u) : z;
    for (int t0 = 0; t0 < seg_chunks; t0 += 32 * kU) {     // warp-uniform trip count (syncwarp inside)
        const int jj0 = t0 + lane;
        uint4 xv[kU], rn[kU];
        {
#pragma unroll
            for (int u = 0; u < kU; u++) {                    // reference samples of the NEXT trip
                const int jn = jj0 + 32 * (u + kU);
                rn[u] = jn < seg_chunks ? __ldg(rin + jn) : z;
            }
#pragma unroll
            for (int u = 0; u < kU; u++) {
                const int jj = jj0 + 32 * u;
                xv[u] = read_shifted<K, TOSIGNED>(sx, jj < seg_chunks ? jj : 0, jj < seg_chunks ? a_base + jj : -2, nchunks);
            }
            __syncwarp();
#pragma unroll
            for (int u = 0; u < kU; u++) {
                const int jj = jj0 + 32 * u;
                if (jj < seg_chunks) sx[jj] = xv[u];
            }
        }
#pragma unroll
        for (int u = 0; u < kU; u++) {
            const uint32_t xw[4] = { xv[u].x, xv[u].y, xv[u].z, xv[u].w }, rw[4] = { rv[u].x, rv[u].y, rv[u].z, rv[u].w };
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t xs = xw[k] ^ TOSIGNED, rs = rw[k] ^ TOSIGNED;
                // bytes [xr0 xi0 xr1 xi1] . [rr0 ri0 rr1 ri1]
                acc[0] = __dp4a((int)xs, (int)rs, acc[0]);
                const uint32_t xsw = __byte_perm(xs, 0, 0x2301);               // [xi0 xr0 xi1 xr1]
                acc[1] = __dp4a((int)xsw, (int)rs, acc[1]);
                acc[2] = __dp4a((int)(xsw & 0xFF00FF00u), (int)rs, acc[2]);
            }
            rv[u] = rn[u];
        }
    }
    __syncwarp();
}

// realignment only (noise source off, or a row longer than the resident segment in pass 2)
template <int K, uint32_t TOSIGNED>
__device__ __forceinline__ void realign_segment(uint4 *sx, int seg_chunks, int a_base, int nchunks, int lane) {
    for (int t0 = 0; t0 < seg_chunks; t0 += 32 * kU) {     // warp-uniform trip count (syncwarp inside)
        const int jj0 = t0 + lane;
        uint4 xv[kU];
#pragma unroll
        for (int u = 0; u < kU; u++) {
            const int jj = jj0 + 32 * u;
            xv[u] = read_shifted<K, TOSIGNED>(sx, jj < seg_chunks ? jj : 0, jj < seg_chunks ? a_base + jj : -2, nchunks);
        }
        __syncwarp();
#pragma unroll
        for (int u = 0; u < kU; u++) {
            const int jj = jj0 + 32 * u;
            if (jj < seg_chunks) sx[jj] = xv[u];
        }
    }
    __syncwarp();
}

// ---- pass 2 over one resident segment: rotate (csdrdevice::phasecorrect), pack (cpacketize::write)
// reads the aligned x' pass 1 left in the buffer
template <uint32_t TORAW, bool CF>
__device__ __forceinline__ void rotate_segment(const uint4 *sx, uint4 *__restrict__ col, float2 *cfrow, int seg_chunks,
                                               int lane, const RotConst &rk) {
    for (int jj0 = lane; jj0 < seg_chunks; jj0 += 32 * kU) {
#pragma unroll
        for (int u = 0; u < kU; u++) {
            const int jj = jj0 + 32 * u;
            if (jj < seg_chunks) {
                const uint4 xv = sx[jj];
                float2 *cf = CF ? cfrow + (size_t)jj * 8 : nullptr;
                uint4 o;
                o.x = rotate_pack_word<CF>(xv.x ^ TORAW, rk, cf);
                o.y = rotate_pack_word<CF>(xv.y ^ TORAW, rk, cf + 2);
                o.z = rotate_pack_word<CF>(xv.z ^ TORAW, rk, cf + 4);
                o.w = rotate_pack_word<CF>(xv.w ^ TORAW, rk, cf + 6);
                col[jj] = o;
            }
        }
    }
}

#define COH_SWITCH_K(K_, CALL)                                                              \
    switch (K_) {                                                                           \
        case 0: { constexpr int KK = 0; CALL; } break;                                      \
        case 1: { constexpr int KK = 1; CALL; } break;                                      \
        case 2: { constexpr int KK = 2; CALL; } break;                                      \
        case 3: { constexpr int KK = 3; CALL; } break;                                      \
        case 4: { constexpr int KK = 4; CALL; } break;                                      \
        case 5: { constexpr int KK = 5; CALL; } break;                                      \
        case 6: { constexpr int KK = 6; CALL; } break;                                      \
        default: { constexpr int KK = 7; CALL; } break;                                     \
    }

__device__ __forceinline__ bool block_enabled(const ApplyLaunch &p, int j) {
    return p.refnoise ? (p.refnoise[j] != 0) : (p.refnoise_scalar != 0);
}

// p <- alpha*q + (1-alpha)*p with individually rounded operations (csdrdevice.cc:67); aq = alpha*q
__device__ __forceinline__ void iir_step(float2 &pp, float2 aq, float na) {
    pp.x = __fadd_rn(aq.x, __fmul_rn(na, pp.x));
    pp.y = __fadd_rn(aq.y, __fmul_rn(na, pp.y));
}

template <bool SIGNED_IN, bool CF>
__global__ void __launch_bounds__(kApplyThreads, 1)
k_apply(const ApplyLaunch p, uint32_t n_items, uint32_t total_warps) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t s_bar[kWarpsPerCta];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr uint32_t kToSigned = SIGNED_IN ? 0u : 0x80808080u;   // input domain -> two's complement
    constexpr uint32_t kToRaw = SIGNED_IN ? 0x80808080u : 0u;      // input domain -> offset binary
    const uint32_t B = p.B, L = B >> 1, nch = p.nch;
    const int nchunks = (int)(B >> 4);
    const int seg_chunks = nchunks < (int)(kSegBytes >> 4) ? nchunks : (int)(kSegBytes >> 4);
    const int nseg = nchunks / seg_chunks;
    const size_t hdr = p.noheader ? 0 : 16 + 4 * (size_t)p.out_nch;
    const float a_w = p.alpha, na_w = 1.0f - p.alpha;
    uint4 *sx = reinterpret_cast<uint4 *>(smem_raw + (size_t)warp * kBufBytes);
    float2 *s_aq = reinterpret_cast<float2 *>(smem_raw + (size_t)warp * kBufBytes + kSegBytes + 32);   // 64 entries
    const uint32_t sx_addr = smem_u32(sx), bar = smem_u32(&s_bar[warp]);
    uint32_t n_fetch = 0;                                  // completed bulk copies -> mbarrier parity

    if (lane == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    // segment s of the row into the warp's buffer: buffer chunk jj <- row chunk seg0 + a0 + jj
    auto fetch = [&](const uint8_t *xrow, int s, int a0) {
        const int a_base = s * seg_chunks + a0;
        const int g0 = a_base > 0 ? a_base : 0;
        const int g1 = a_base + seg_chunks + 1 < nchunks ? a_base + seg_chunks + 1 : nchunks;
        // every lane is done with the old content, and its generic-proxy writes to the buffer
        // (pass 1 writes x' back) are ordered before the asynchronous-proxy write of the copy
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (g1 > g0) {
            if (lane == 0)
                tma_row_load(sx_addr + (uint32_t)(g0 - a_base) * 16u, xrow + (size_t)g0 * 16, (uint32_t)(g1 - g0) * 16u, bar);
            mbar_wait(bar, n_fetch & 1u);
            n_fetch++;
        }
        return a_base;
    };

    uint32_t item = 0;
    if (lane == 0) item = atomicAdd(p.ticket, 1u);
    item = __shfl_sync(0xffffffffu, item, 0);

    while (item < n_items) {
        const uint32_t bl = item / nch, c = item - bl * nch, b = p.block0 + bl;
        const size_t bc = (size_t)b * nch + c;
        const uint8_t *xrow = p.iq + bc * B;
        const uint4 *rin = reinterpret_cast<const uint4 *>(p.iq + (size_t)b * nch * B);
        int8_t *pkt = (b < p.out_skip ? p.packets_skip + (size_t)b * p.pitch
                                      : p.packets + (size_t)(p.out_block0 + b - p.out_skip) * p.pitch) + p.lead;
        uint4 *col = reinterpret_cast<uint4 *>(pkt + hdr + (size_t)(c ? p.out_col0 + c - 1 : 0) * B);

        if (c == 0) {
            // reference column: raw signed bytes (cpacketize::write(0,..,int8*), cpacketizer.cc:137-156)
            if (p.write_ref || CF) {
                // four independent 16-byte loads in flight per lane (a serial copy of the 16 KiB row would hold
                // this warp for ~30 us of load latency)
                for (int j0 = lane; j0 < nchunks; j0 += 128) {
                    uint4 v[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) if (j0 + 32 * u < nchunks) v[u] = __ldg(rin + j0 + 32 * u);
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const int j = j0 + 32 * u;
                        if (j >= nchunks) break;
                        v[u].x ^= kToSigned; v[u].y ^= kToSigned; v[u].z ^= kToSigned; v[u].w ^= kToSigned;
                        if (p.write_ref) col[j] = v[u];
                        if (CF) {
                            float2 *o = p.cf32 + bc * L + (size_t)j * 8;
                            const uint32_t wv[4] = { v[u].x, v[u].y, v[u].z, v[u].w };
                            const float cc = 0x1.020408p-7f;
#pragma unroll
                            for (int k = 0; k < 4; k++) {
                                int w = (int)wv[k];
                                o[2 * k] = make_float2(__fmul_rn((float)(signed char)(w & 0xff), cc), __fmul_rn((float)(signed char)((w >> 8) & 0xff), cc));
                                o[2 * k + 1] = make_float2(__fmul_rn((float)(signed char)((w >> 16) & 0xff), cc), __fmul_rn((float)(signed char)((w >> 24) & 0xff), cc));
                            }
                        }
                    }
                }
            }
            if (!p.noheader && p.write_hdr) {
                // struct hdr0 (include/cpacketizer.h:32-37) + per-channel readcnt (cpacketizer.cc:142)
                uint32_t *h32 = reinterpret_cast<uint32_t *>(pkt);
                if (lane == 0) { h32[0] = p.globalseqn0 + b; h32[1] = p.out_nch; h32[2] = L; h32[3] = 0; }
                for (uint32_t k = lane; k < p.out_nch; k += 32)
                    h32[4 + k] = p.readcnt ? p.readcnt[(size_t)b * p.out_nch + k] : p.readcnt0 + b;
            }
            if (lane == 0) {
                if (p.lag) p.lag[bc] = 0.0f;
                if (p.mag) p.mag[bc] = 0.0f;
                if (p.phase) p.phase[bc] = make_float2(1.0f, 0.0f);
                if (p.measured) p.measured[bc] = 0;
                if (b == p.n_blocks - 1) p.state_out[0] = p.state_in[0];
            }
        } else {
            // ---- lag that applies to this block (set_lag, include/csdrdevice.h:138-151) -----------
            const int32_t src = p.src_block ? p.src_block[bc] : (p.lag_all ? (int32_t)b : -1);
            int32_t lagi; float lagf, magf;
            if (src >= 0) {
                const PeakRec pk = p.peaks[(size_t)src * nch + c];
                lagi = (int32_t)pk.idx - (int32_t)L;
                lagf = (float)pk.idx - (float)L;                              // ccoherent.cc:218-232
                magf = __fsqrt_rn(__fdiv_rn(pk.val, (float)L));               // ccoherent.cc:203-204
            } else {
                const ChanState st = p.state_in[c];
                lagi = st.held_lag; lagf = st.lag; magf = st.mag;
            }
            const int sh = p.apply_shift ? lagi : 0;
            const int a0 = (2 * sh) >> 4, K = ((2 * sh) & 15) >> 1;          // arithmetic shift = floor
            const bool enabled = block_enabled(p, (int)b);

            // ---- pass 1: raw phase estimate q = conj(S)/|S| --------------------------------------
            int a_base = 0;
            float2 q = make_float2(1.0f, 0.0f);
            if (enabled || nseg == 1) a_base = fetch(xrow, 0, a0);
            if (!enabled && nseg == 1) {
                COH_SWITCH_K(K, (realign_segment<KK, kToSigned>(sx, seg_chunks, a_base, nchunks, lane)));
            }
            if (enabled) {
                int acc[3] = { 0, 0, 0 };
                for (int s = 0; s < nseg; s++) {
                    if (s) a_base = fetch(xrow, s, a0);
                    COH_SWITCH_K(K, (dot_segment<KK, kToSigned>(sx, rin + (size_t)s * seg_chunks, seg_chunks, a_base, nchunks, lane, acc)));
                }
                long long sre, sim;
                if (L <= 32768) {                     // |S| < 2^31: warp sums fit 32 bits
                    sre = __reduce_add_sync(0xffffffffu, acc[0]);
                    sim = __reduce_add_sync(0xffffffffu, acc[1] - 2 * acc[2]);         // sum xi*rr - xr*ri
                } else {
                    sre = acc[0]; sim = (long long)acc[1] - 2ll * acc[2];
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1) {
                        sre += __shfl_xor_sync(0xffffffffu, sre, off);
                        sim += __shfl_xor_sync(0xffffffffu, sim, off);
                    }
                }
                // csdrdevice.cc:63, individually rounded operations
                const float cr = (float)sre, ci = (float)sim;
                const float m = __fsqrt_rn(__fmaf_rn(cr, cr, __fmul_rn(ci, ci)));
                const float inv = __fdiv_rn(1.0f, m);
                q = make_float2(__fmul_rn(cr, inv), -__fmul_rn(ci, inv));
                if (lane == 0) publish16(&p.q[bc], __float_as_uint(q.x), __float_as_uint(q.y), p.epoch);
            }

            // ---- decoupled look-back over the previous enabled blocks of this channel -------------
            // window [jstart, b): the newest `lookback` enabled blocks, or everything back to the
            // start of the batch (then the carried state is the exact initial value)
            const int D = p.lookback;
            int jstart; bool have;
            if (p.refnoise) {
                int cnt = 0; jstart = (int)b;
                while (jstart > 0 && cnt < D) {
                    const int j = jstart - 1 - lane;                         // lane 0 = newest
                    const uint32_t m = __ballot_sync(0xffffffffu, j >= 0 && p.refnoise[j] != 0);
                    const int pc = __popc(m);
                    if (cnt + pc >= D) { jstart = jstart - 1 - (int)__fns(m, 0, D - cnt); cnt = D; break; }
                    cnt += pc; jstart = jstart > 32 ? jstart - 32 : 0;
                }
                have = cnt < D;
                if (have) jstart = 0;
            } else if (p.refnoise_scalar) {
                have = (int)b < D; jstart = have ? 0 : (int)b - D;
            } else { have = true; jstart = (int)b; }
            float2 pp = make_float2(0.0f, 0.0f);
            if (have) pp = p.state_in[c].phasecorrprev;
            for (int base = jstart; base < (int)b; base += 32) {              // oldest first
                const int j = base + lane;
                const bool en = j < (int)b && block_enabled(p, j);
                float2 qv = make_float2(0.0f, 0.0f);
                if (en) {
                    const uint4 v = poll16(&p.q[(size_t)j * nch + c], p.epoch);
                    qv = make_float2(__uint_as_float(v.x), __uint_as_float(v.y));
                }
                const uint32_t m = __ballot_sync(0xffffffffu, en);
                const int pos = __popc(m & ((1u << lane) - 1u)), cnt = __popc(m);
                __syncwarp();
                if (en) s_aq[pos] = make_float2(__fmul_rn(a_w, qv.x), __fmul_rn(a_w, qv.y));
                int i0 = 0;
                if (!have && m) {                       // truncated tail: p := oldest estimate of the window
                    const int f = __ffs(m) - 1;
                    pp.x = __shfl_sync(0xffffffffu, qv.x, f); pp.y = __shfl_sync(0xffffffffu, qv.y, f);
                    have = true; i0 = 1;
                }
                __syncwarp();
                // csdrdevice.cc:67, each operation rounded on its own.  Scalar intrinsics on purpose:
                // ptxas 12.9 contracts mul.rn.f32x2 + add.rn.f32x2 into one FFMA2, which changes the
                // last bit whenever (1-alpha) is not a power of two.
#pragma unroll 4
                for (int i = i0; i < cnt; i++) {
                    const float2 aq = s_aq[i];
                    pp.x = __fadd_rn(aq.x, __fmul_rn(na_w, pp.x));
                    pp.y = __fadd_rn(aq.y, __fmul_rn(na_w, pp.y));
                }
            }
            if (enabled) iir_step(pp, make_float2(__fmul_rn(a_w, q.x), __fmul_rn(a_w, q.y)), na_w);   // csdrdevice.cc:67
            if (lane == 0) {
                if (p.phase) p.phase[bc] = pp;
                if (p.lag) p.lag[bc] = lagf;
                if (p.mag) p.mag[bc] = magf;
                if (p.measured) p.measured[bc] = (src == (int32_t)b) ? 1 : 0;
                if (b == p.n_blocks - 1) {
                    ChanState so; so.phasecorr = pp; so.phasecorrprev = pp;   // csdrdevice.cc:68
                    so.lag = lagf; so.mag = magf; so.held_lag = lagi; so.pad = 0;
                    p.state_out[c] = so;
                }
            }

            // ---- pass 2: rotate and pack, from the bytes that are still resident -------------------
            RotConst rk;
            rk.px = make_float2(pp.x, pp.x); rk.py = make_float2(pp.y, pp.y); rk.npy = make_float2(-pp.y, -pp.y);
            for (int s = 0; s < nseg; s++) {
                if (nseg > 1) {                                               // long rows: again, from L2
                    a_base = fetch(xrow, s, a0);
                    COH_SWITCH_K(K, (realign_segment<KK, kToSigned>(sx, seg_chunks, a_base, nchunks, lane)));
                }
                float2 *cfrow = CF ? p.cf32 + bc * L + (size_t)s * seg_chunks * 8 : nullptr;
                rotate_segment<kToRaw, CF>(sx, col + (size_t)s * seg_chunks, cfrow, seg_chunks, lane, rk);
            }
        }
        // the next ticket is drawn only now: a ticket held while another item is still in its
        // look-back would make every waiter on it wait for that look-back too (convoy)
        if (lane == 0) item = atomicAdd(p.ticket, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
    }
    // every warp ends on exactly one ticket >= n_items; the largest one is the last ever drawn
    if (lane == 0 && item == n_items + total_warps - 1) atomicExch(p.ticket, 0u);
}

template <bool SIGNED_IN, bool CF>
static cudaError_t launch_apply_t(const ApplyLaunch &p, uint32_t n_items, cudaStream_t st) {
    const size_t smem = (size_t)kWarpsPerCta * kBufBytes;
    static bool attr_done[64] = {};
    int dev = 0; cudaGetDevice(&dev);
    if (!attr_done[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(k_apply<SIGNED_IN, CF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        attr_done[dev & 63] = true;
    }
    int occ = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_apply<SIGNED_IN, CF>, kApplyThreads, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    uint32_t grid = (n_items + kWarpsPerCta - 1) / kWarpsPerCta;
    uint32_t cap = (uint32_t)(p.sm_count > 0 ? p.sm_count : 148) * (uint32_t)occ;
    if (p.max_ctas && p.max_ctas < cap) cap = p.max_ctas;
    if (grid > cap) grid = cap;
    k_apply<SIGNED_IN, CF><<<grid, kApplyThreads, smem, st>>>(p, n_items, grid * kWarpsPerCta);
    return cudaGetLastError();
}

cudaError_t launch_apply(const ApplyLaunch &p, cudaStream_t st) {
    const uint32_t n_items = p.n_launch_blocks * p.nch;
    if (n_items == 0) return cudaSuccess;
    if (p.lookback < 1 || p.lookback > kMaxLookback) return cudaErrorInvalidValue;
    const bool sgn = p.xormask == 0, cf = p.cf32 != nullptr;
    if (sgn && cf) return launch_apply_t<true, true>(p, n_items, st);
    if (sgn)       return launch_apply_t<true, false>(p, n_items, st);
    if (cf)        return launch_apply_t<false, true>(p, n_items, st);
    return launch_apply_t<false, false>(p, n_items, st);
}

}  // namespace coh
