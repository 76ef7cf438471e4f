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
// One CTA handles one (block, channel, 16 KiB part) item; items are handed out through an
// atomic ticket so that every item with a smaller ticket is already running -- which makes the
// waits below (sibling parts, look-back over earlier blocks) deadlock-free.
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

namespace coh {

constexpr int kApplyThreads = 256;
constexpr int kChunksPerThread = 4;                       // 4 x 16 B = 64 B of x per thread
constexpr uint32_t kPartBytes = kApplyThreads * kChunksPerThread * 16;   // 16 KiB
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

// ---- shifted load: 16 bytes of x'[n] = x[n + sh] (raw domain, zero_code outside the block) ---
// j = 16-byte chunk index inside the channel-block.  The byte offset 16*j + 2*sh has the same
// residue mod 16 for every chunk, so the word select below is uniform across the CTA.
__device__ __forceinline__ uint4 load_x_chunk(const uint4 *x4, int j, int sh, int nchunks, uint32_t zero_code) {
    const uint4 z = make_uint4(zero_code, zero_code, zero_code, zero_code);
    const int o = 16 * j + 2 * sh;                       // may be negative
    const int a = o >> 4, rem = o & 15;                  // arithmetic shift = floor
    uint4 A = (a >= 0 && a < nchunks) ? __ldg(x4 + a) : z;
    if (rem == 0) return A;
    uint4 Bv = (a + 1 >= 0 && a + 1 < nchunks) ? __ldg(x4 + a + 1) : z;
    uint32_t w0, w1, w2, w3, w4;
    switch (rem >> 2) {
        case 0:  w0 = A.x; w1 = A.y; w2 = A.z; w3 = A.w; w4 = Bv.x; break;
        case 1:  w0 = A.y; w1 = A.z; w2 = A.w; w3 = Bv.x; w4 = Bv.y; break;
        case 2:  w0 = A.z; w1 = A.w; w2 = Bv.x; w3 = Bv.y; w4 = Bv.z; break;
        default: w0 = A.w; w1 = Bv.x; w2 = Bv.y; w3 = Bv.z; w4 = Bv.w; break;
    }
    if (rem & 2)
        return make_uint4(__funnelshift_r(w0, w1, 16), __funnelshift_r(w1, w2, 16),
                          __funnelshift_r(w2, w3, 16), __funnelshift_r(w3, w4, 16));
    return make_uint4(w0, w1, w2, w3);
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

template <bool SIGNED_IN, bool CF>
__global__ void __launch_bounds__(kApplyThreads)
k_apply(const ApplyLaunch p, uint32_t ticket0, uint32_t parts) {
    __shared__ uint32_t s_item;
    __shared__ int s_red[3][kApplyThreads / 32];
    __shared__ float2 s_p;
    extern __shared__ float2 s_q[];                       // lookback + 32 entries

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_item = atomicAdd(p.ticket, 1u) - ticket0;
    __syncthreads();
    const uint32_t item = s_item;
    const uint32_t part = item % parts;
    const uint32_t c = (item / parts) % p.nch;
    const uint32_t b = p.block0 + item / (parts * p.nch);
    const uint32_t B = p.B, L = B >> 1;
    const uint32_t part_bytes = B < kPartBytes ? B : kPartBytes;
    const int nchunk_part = part_bytes >> 4, nchunks = B >> 4;
    const int chunk0 = part * (kPartBytes >> 4);
    const size_t bc = (size_t)b * p.nch + c;
    // input bytes -> signed / offset-binary views: one of the two is the raw word
    constexpr uint32_t kToSigned = SIGNED_IN ? 0u : 0x80808080u;

    const uint4 *xin = reinterpret_cast<const uint4 *>(p.iq + bc * B);
    const uint4 *rin = reinterpret_cast<const uint4 *>(p.iq + (size_t)b * p.nch * B);
    const size_t hdr = p.noheader ? 0 : 16 + 4 * (size_t)p.nch;
    int8_t *pkt = p.packets + (size_t)b * p.pitch + p.lead;
    uint4 *col = reinterpret_cast<uint4 *>(pkt + hdr + (size_t)c * B);

    if (c == 0) {
        // reference column: raw signed bytes (cpacketize::write(0,..,int8*), cpacketizer.cc:137-156)
#pragma unroll
        for (int i = 0; i < kChunksPerThread; i++) {
            int jl = tid + i * kApplyThreads;
            if (jl < nchunk_part) {
                uint4 v = __ldg(rin + chunk0 + jl);
                v.x ^= kToSigned; v.y ^= kToSigned; v.z ^= kToSigned; v.w ^= kToSigned;
                col[chunk0 + jl] = v;
                if (CF) {
                    float2 *o = p.cf32 + bc * L + (size_t)(chunk0 + jl) * 8;
                    const uint32_t wv[4] = { v.x, v.y, v.z, v.w };
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
        if (part == 0) {
            if (!p.noheader) {
                // struct hdr0 (include/cpacketizer.h:32-37) + per-channel readcnt (cpacketizer.cc:142)
                uint32_t *h32 = reinterpret_cast<uint32_t *>(pkt);
                if (tid == 0) { h32[0] = p.globalseqn0 + b; h32[1] = p.nch; h32[2] = L; h32[3] = 0; }
                for (uint32_t k = tid; k < p.nch; k += kApplyThreads)
                    h32[4 + k] = p.readcnt ? p.readcnt[(size_t)b * p.nch + k] : p.readcnt0 + b;
            }
            if (tid == 0) {
                if (p.lag) p.lag[bc] = 0.0f;
                if (p.mag) p.mag[bc] = 0.0f;
                if (p.phase) p.phase[bc] = make_float2(1.0f, 0.0f);
                if (p.measured) p.measured[bc] = 0;
                if (b == p.n_blocks - 1) p.state_out[0] = p.state_in[0];
            }
        }
        return;
    }

    // ---- lag that applies to this block (set_lag, include/csdrdevice.h:138-151) ---------------
    int32_t src = p.src_block ? p.src_block[bc] : (p.lag_all ? (int32_t)b : -1);
    int32_t lagi; float lagf, magf;
    if (src >= 0) {
        PeakRec pk = p.peaks[(size_t)src * p.nch + c];
        lagi = (int32_t)pk.idx - (int32_t)L;
        lagf = (float)pk.idx - (float)L;                              // ccoherent.cc:218-232
        magf = __fsqrt_rn(__fdiv_rn(pk.val, (float)L));               // ccoherent.cc:203-204
    } else {
        ChanState st = p.state_in[c];
        lagi = st.held_lag; lagf = st.lag; magf = st.mag;
    }
    const int sh = p.apply_shift ? lagi : 0;

    // ---- load x' (kept in registers) and r, exact integer conjugate dot product ---------------
    uint4 xv[kChunksPerThread];                           // signed bytes
    int acc_re = 0, acc_u = 0, acc_v = 0;
#pragma unroll
    for (int i = 0; i < kChunksPerThread; i++) {
        int jl = tid + i * kApplyThreads;
        uint4 xr = make_uint4(0, 0, 0, 0), rr = make_uint4(0, 0, 0, 0);
        if (jl < nchunk_part) {
            xr = load_x_chunk(xin, chunk0 + jl, sh, nchunks, kToSigned);
            rr = __ldg(rin + chunk0 + jl);
            xr.x ^= kToSigned; xr.y ^= kToSigned; xr.z ^= kToSigned; xr.w ^= kToSigned;
            rr.x ^= kToSigned; rr.y ^= kToSigned; rr.z ^= kToSigned; rr.w ^= kToSigned;
        }
        xv[i] = xr;
        const uint32_t xw[4] = { xr.x, xr.y, xr.z, xr.w }, rw[4] = { rr.x, rr.y, rr.z, rr.w };
#pragma unroll
        for (int k = 0; k < 4; k++) {
            // bytes [xr0 xi0 xr1 xi1] . [rr0 ri0 rr1 ri1]
            acc_re = __dp4a((int)xw[k], (int)rw[k], acc_re);                   // sum xr*rr + xi*ri
            uint32_t xsw = __byte_perm(xw[k], 0, 0x2301);                      // [xi0 xr0 xi1 xr1]
            acc_u = __dp4a((int)xsw, (int)rw[k], acc_u);                       // sum xi*rr + xr*ri
            acc_v = __dp4a((int)(xsw & 0xFF00FF00u), (int)rw[k], acc_v);       // sum xr*ri
        }
    }
    acc_re = __reduce_add_sync(0xffffffffu, acc_re);
    acc_u = __reduce_add_sync(0xffffffffu, acc_u);
    acc_v = __reduce_add_sync(0xffffffffu, acc_v);
    if (lane == 0) { s_red[0][warp] = acc_re; s_red[1][warp] = acc_u; s_red[2][warp] = acc_v; }
    __syncthreads();

    const bool enabled = p.refnoise ? (p.refnoise[b] != 0) : (p.refnoise_scalar != 0);
    if (warp == 0) {
        long long sre = 0, sim = 0;
        if (enabled) {
            int vre = lane < kApplyThreads / 32 ? s_red[0][lane] : 0;
            int vu = lane < kApplyThreads / 32 ? s_red[1][lane] : 0;
            int vv = lane < kApplyThreads / 32 ? s_red[2][lane] : 0;
            vre = __reduce_add_sync(0xffffffffu, vre);
            vu = __reduce_add_sync(0xffffffffu, vu);
            vv = __reduce_add_sync(0xffffffffu, vv);
            int pre = vre, pim = vu - 2 * vv;                                  // sum xi*rr - xr*ri
            if (parts > 1) {
                // exchange the partial sums of the sibling parts (tickets are adjacent)
                if (lane == 0) publish16(&p.partial[bc * parts + part], (uint32_t)pre, (uint32_t)pim, p.epoch);
                for (uint32_t k = lane; k < parts; k += 32) {
                    uint4 v = poll16(&p.partial[bc * parts + k], p.epoch);
                    sre += (int)v.x; sim += (int)v.y;
                }
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) {
                    sre += __shfl_xor_sync(0xffffffffu, sre, off);
                    sim += __shfl_xor_sync(0xffffffffu, sim, off);
                }
            } else { sre = pre; sim = pim; }
        }
        // raw estimate q = conj(S)/|S| (csdrdevice.cc:63), individually rounded operations
        float2 q = make_float2(1.0f, 0.0f);
        if (enabled) {
            float cr = (float)sre, ci = (float)sim;
            float m = __fsqrt_rn(__fmaf_rn(cr, cr, __fmul_rn(ci, ci)));
            float inv = __fdiv_rn(1.0f, m);
            q = make_float2(__fmul_rn(cr, inv), -__fmul_rn(ci, inv));
            if (lane == 0 && part == 0) publish16(&p.q[bc], __float_as_uint(q.x), __float_as_uint(q.y), p.epoch);
        }
        // ---- decoupled look-back over earlier blocks of this channel --------------------------
        const int D = p.lookback;
        int cnt = 0;
        for (int jhi = (int)b - 1; jhi >= 0 && cnt < D; jhi -= 32) {
            const int j = jhi - lane;
            const bool en = j >= 0 && (p.refnoise ? (p.refnoise[j] != 0) : (p.refnoise_scalar != 0));
            float2 qv = make_float2(0.f, 0.f);
            if (en) {
                uint4 v = poll16(&p.q[(size_t)j * p.nch + c], p.epoch);
                qv = make_float2(__uint_as_float(v.x), __uint_as_float(v.y));
            }
            const uint32_t m = __ballot_sync(0xffffffffu, en);
            const int pos = cnt + __popc(m & ((1u << lane) - 1u));
            if (en) s_q[pos] = qv;                                            // newest first
            cnt += __popc(m);
        }
        __syncwarp();
        if (lane == 0) {
            const float a = p.alpha, na = 1.0f - p.alpha;
            float2 pp; int i;
            if (cnt < D) { pp = p.state_in[c].phasecorrprev; i = cnt - 1; }   // reached batch start: exact
            else { pp = s_q[D - 1]; i = D - 2; }                              // truncated tail
#pragma unroll 4
            for (; i >= 0; i--) {
                float2 qq = s_q[i];
                pp.x = __fadd_rn(__fmul_rn(a, qq.x), __fmul_rn(na, pp.x));    // csdrdevice.cc:67
                pp.y = __fadd_rn(__fmul_rn(a, qq.y), __fmul_rn(na, pp.y));
            }
            if (enabled) {
                pp.x = __fadd_rn(__fmul_rn(a, q.x), __fmul_rn(na, pp.x));
                pp.y = __fadd_rn(__fmul_rn(a, q.y), __fmul_rn(na, pp.y));
            }
            s_p = pp;
            if (part == 0) {
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
        }
    }
    __syncthreads();
    const float2 pc = s_p;
    RotConst rk;
    rk.px = make_float2(pc.x, pc.x); rk.py = make_float2(pc.y, pc.y); rk.npy = make_float2(-pc.y, -pc.y);

    // ---- rotate (csdrdevice::phasecorrect) and pack (cpacketize::write) ------------------------
#pragma unroll
    for (int i = 0; i < kChunksPerThread; i++) {
        int jl = tid + i * kApplyThreads;
        if (jl < nchunk_part) {
            float2 *cf = CF ? p.cf32 + bc * L + (size_t)(chunk0 + jl) * 8 : nullptr;
            uint4 o;
            o.x = rotate_pack_word<CF>(xv[i].x ^ 0x80808080u, rk, cf);
            o.y = rotate_pack_word<CF>(xv[i].y ^ 0x80808080u, rk, cf + 2);
            o.z = rotate_pack_word<CF>(xv[i].z ^ 0x80808080u, rk, cf + 4);
            o.w = rotate_pack_word<CF>(xv[i].w ^ 0x80808080u, rk, cf + 6);
            col[chunk0 + jl] = o;
        }
    }
}

uint32_t apply_parts(uint32_t B) { return B > kPartBytes ? B / kPartBytes : 1; }

cudaError_t launch_apply(const ApplyLaunch &p, cudaStream_t st) {
    const uint32_t parts = apply_parts(p.B);
    const uint32_t n_items = p.n_launch_blocks * p.nch * parts;
    if (n_items == 0) return cudaSuccess;
    if (p.lookback < 1 || p.lookback > kMaxLookback) return cudaErrorInvalidValue;
    const size_t smem = (size_t)(p.lookback + 32) * sizeof(float2);
    const bool sgn = p.xormask == 0, cf = p.cf32 != nullptr;
    if (sgn && cf)       k_apply<true, true><<<n_items, kApplyThreads, smem, st>>>(p, p.ticket0, parts);
    else if (sgn)        k_apply<true, false><<<n_items, kApplyThreads, smem, st>>>(p, p.ticket0, parts);
    else if (cf)         k_apply<false, true><<<n_items, kApplyThreads, smem, st>>>(p, p.ticket0, parts);
    else                 k_apply<false, false><<<n_items, kApplyThreads, smem, st>>>(p, p.ticket0, parts);
    return cudaGetLastError();
}

}  // namespace coh
