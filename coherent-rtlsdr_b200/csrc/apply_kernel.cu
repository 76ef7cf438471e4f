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
#include "kernels.h"

namespace coh {

constexpr int kApplyThreads = 256;
constexpr int kChunksPerThread = 4;                       // 4 x 16 B = 64 B of x per thread
constexpr uint32_t kPartBytes = kApplyThreads * kChunksPerThread * 16;   // 16 KiB
constexpr int kMaxLookback = 2048;

__device__ __forceinline__ uint32_t ld_acquire(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(uint32_t *p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// 16 bytes of x'[n] = x[n + sh] (raw domain, zero_code outside the block); j = 16-byte chunk
// index inside the channel-block
__device__ __forceinline__ uint4 load_x_chunk(const uint8_t *x, int j, int sh, uint32_t B, uint32_t zero_code) {
    uint4 r;
    if ((sh & 7) == 0) {
        int jj = j + (sh >> 3);
        if (jj >= 0 && jj < (int)(B >> 4)) r = __ldg(reinterpret_cast<const uint4 *>(x) + jj);
        else r = make_uint4(zero_code, zero_code, zero_code, zero_code);
    } else {
        const int o = 16 * j + 2 * sh;                   // byte offset, may be negative
        const int w0 = o >> 2, sub = o & 3;              // arithmetic shift = floor
        uint32_t w[5];
#pragma unroll
        for (int i = 0; i < 5; i++) {
            int idx = w0 + i;
            w[i] = (idx >= 0 && idx < (int)(B >> 2)) ? __ldg(reinterpret_cast<const uint32_t *>(x) + idx) : zero_code;
        }
        if (sub) {
            r.x = __funnelshift_r(w[0], w[1], 16); r.y = __funnelshift_r(w[1], w[2], 16);
            r.z = __funnelshift_r(w[2], w[3], 16); r.w = __funnelshift_r(w[3], w[4], 16);
        } else { r.x = w[0]; r.y = w[1]; r.z = w[2]; r.w = w[3]; }
    }
    return r;
}

// exact (float)int8 * float(1/127) of byte k of the offset-binary word xu
// (cdsp::convtofloat, src/cdsp.cc:41-44: volk multiplies by the rounded reciprocal)
template <int K>
__device__ __forceinline__ float byte_to_float(uint32_t xu) {
    float f = __uint_as_float(__byte_perm(xu, 0x4B000000u, 0x7440 | K));     // 2^23 + u8
    return __fmul_rn(f - 8388736.0f, 0x1.020408p-7f);                          // (u8 - 128) / 127
}

// volk_32f_s32f_convert_8i (src/cdsp.cc:51-54): clamp(v*127) to [-128,127], round half even
__device__ __forceinline__ uint32_t float_to_s8_bits(float y) {
    float v = __fmul_rn(y, 127.0f);
    v = fminf(fmaxf(v, -128.0f), 127.0f);
    return __float_as_uint(__fadd_rn(v, 12582912.0f));   // low byte = rint(v) two's complement
}

// rotate two samples (one 32-bit word of signed bytes) by p and pack back to signed bytes
__device__ __forceinline__ uint32_t rotate_pack_word(uint32_t xs, float2 p, float2 *cf_out) {
    const uint32_t xu = xs ^ 0x80808080u;
    float r0 = byte_to_float<0>(xu), i0 = byte_to_float<1>(xu);
    float r1 = byte_to_float<2>(xu), i1 = byte_to_float<3>(xu);
    // volk_32fc_s32fc_multiply_32fc, x86 AVX+FMA kernel rounding: fmaddsub(a, re(p), swap(a)*im(p))
    float yr0 = __fmaf_rn(r0, p.x, -__fmul_rn(i0, p.y)), yi0 = __fmaf_rn(i0, p.x, __fmul_rn(r0, p.y));
    float yr1 = __fmaf_rn(r1, p.x, -__fmul_rn(i1, p.y)), yi1 = __fmaf_rn(i1, p.x, __fmul_rn(r1, p.y));
    if (cf_out) { cf_out[0] = make_float2(yr0, yi0); cf_out[1] = make_float2(yr1, yi1); }
    uint32_t b0 = float_to_s8_bits(yr0), b1 = float_to_s8_bits(yi0);
    uint32_t b2 = float_to_s8_bits(yr1), b3 = float_to_s8_bits(yi1);
    uint32_t lo = __byte_perm(b0, b1, 0x0040), hi = __byte_perm(b2, b3, 0x0040);
    return __byte_perm(lo, hi, 0x5410);
}

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
    const int nchunk_part = part_bytes >> 4;
    const int chunk0 = part * (kPartBytes >> 4);
    const size_t bc = (size_t)b * p.nch + c;

    const uint8_t *xin = p.iq + bc * B;
    const uint8_t *rin = p.iq + (size_t)b * p.nch * B;
    const size_t hdr = p.noheader ? 0 : 16 + 4 * (size_t)p.nch;
    int8_t *pkt = p.packets + (size_t)b * p.pitch + p.lead;
    uint4 *col = reinterpret_cast<uint4 *>(pkt + hdr + (size_t)c * B);

    if (c == 0) {
        // reference column: raw signed bytes (cpacketize::write(0,..,int8*), cpacketizer.cc:137-156)
#pragma unroll
        for (int i = 0; i < kChunksPerThread; i++) {
            int jl = tid + i * kApplyThreads;
            if (jl < nchunk_part) {
                uint4 v = __ldg(reinterpret_cast<const uint4 *>(rin) + chunk0 + jl);
                v.x ^= p.xormask; v.y ^= p.xormask; v.z ^= p.xormask; v.w ^= p.xormask;
                col[chunk0 + jl] = v;
                if (p.cf32) {
                    float2 *o = p.cf32 + bc * L + (size_t)(chunk0 + jl) * 8;
                    const uint32_t wv[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        uint32_t xu = wv[k] ^ 0x80808080u;
                        o[2 * k] = make_float2(byte_to_float<0>(xu), byte_to_float<1>(xu));
                        o[2 * k + 1] = make_float2(byte_to_float<2>(xu), byte_to_float<3>(xu));
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
    uint4 xv[kChunksPerThread];
    int acc_re = 0, acc_u = 0, acc_v = 0;
#pragma unroll
    for (int i = 0; i < kChunksPerThread; i++) {
        int jl = tid + i * kApplyThreads;
        uint4 xr = make_uint4(0, 0, 0, 0), rr = make_uint4(0, 0, 0, 0);
        if (jl < nchunk_part) {
            xr = load_x_chunk(xin, chunk0 + jl, sh, B, p.xormask);
            rr = __ldg(reinterpret_cast<const uint4 *>(rin) + chunk0 + jl);
            xr.x ^= p.xormask; xr.y ^= p.xormask; xr.z ^= p.xormask; xr.w ^= p.xormask;
            rr.x ^= p.xormask; rr.y ^= p.xormask; rr.z ^= p.xormask; rr.w ^= p.xormask;
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
                if (lane == 0) {
                    p.partial[bc * parts + part] = make_int2(pre, pim);
                    __threadfence();
                    st_release(&p.pflag[bc * parts + part], p.epoch);
                }
                for (uint32_t k = lane; k < parts; k += 32) {
                    while (ld_acquire(&p.pflag[bc * parts + k]) != p.epoch) __nanosleep(64);
                    int2 v = __ldcg(&p.partial[bc * parts + k]);
                    sre += v.x; sim += v.y;
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
            if (lane == 0 && part == 0) {
                p.q[bc] = q;
                __threadfence();
                st_release(&p.flag[bc], p.epoch);
            }
        }
        // ---- decoupled look-back over earlier blocks of this channel --------------------------
        const int D = p.lookback;
        int cnt = 0;
        for (int jhi = (int)b - 1; jhi >= 0 && cnt < D; jhi -= 32) {
            const int j = jhi - lane;
            const bool en = j >= 0 && (p.refnoise ? (p.refnoise[j] != 0) : (p.refnoise_scalar != 0));
            float2 qv = make_float2(0.f, 0.f);
            if (en) {
                const size_t jc = (size_t)j * p.nch + c;
                while (ld_acquire(&p.flag[jc]) != p.epoch) __nanosleep(64);
                qv = __ldcg(&p.q[jc]);
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
    const bool pnan = (pc.x != pc.x) || (pc.y != pc.y);

    // ---- rotate (csdrdevice::phasecorrect) and pack (cpacketize::write) ------------------------
#pragma unroll
    for (int i = 0; i < kChunksPerThread; i++) {
        int jl = tid + i * kApplyThreads;
        if (jl < nchunk_part) {
            float2 *cf = p.cf32 ? p.cf32 + bc * L + (size_t)(chunk0 + jl) * 8 : nullptr;
            uint4 o;
            if (pnan) {      // |S| = 0: the reference propagates NaN (csdrdevice.cc:63); bytes become 0
                o = make_uint4(0, 0, 0, 0);
                if (cf) for (int k = 0; k < 8; k++) cf[k] = make_float2(pc.x, pc.y);
            } else {
                o.x = rotate_pack_word(xv[i].x, pc, cf);
                o.y = rotate_pack_word(xv[i].y, pc, cf ? cf + 2 : nullptr);
                o.z = rotate_pack_word(xv[i].z, pc, cf ? cf + 4 : nullptr);
                o.w = rotate_pack_word(xv[i].w, pc, cf ? cf + 6 : nullptr);
            }
            col[chunk0 + jl] = o;
        }
    }
}

cudaError_t launch_apply(const ApplyLaunch &p, cudaStream_t st) {
    const uint32_t parts = apply_parts(p.B);
    const uint32_t n_items = p.n_launch_blocks * p.nch * parts;
    if (n_items == 0) return cudaSuccess;
    if (p.lookback < 1 || p.lookback > kMaxLookback) return cudaErrorInvalidValue;
    const size_t smem = (size_t)(p.lookback + 32) * sizeof(float2);
    k_apply<<<n_items, kApplyThreads, smem, st>>>(p, p.ticket0, parts);
    return cudaGetLastError();
}

uint32_t apply_parts(uint32_t B) { return B > kPartBytes ? B / kPartBytes : 1; }

}  // namespace coh
