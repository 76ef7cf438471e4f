/*
 * cdsp_oracle.c -- CPU restatement of the coherent-rtlsdr block sync + correction path.
 * TEST INFRASTRUCTURE ONLY (see cdsp_oracle.h).  Build with -ffp-contract=off: every float
 * expression below is meant to round exactly where it is written.
 * Citations are file:line in /root/reference.
 */
#include "cdsp_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

/* ---------------------------------------------------------------- cdsp primitives ------- */

/* src/cdsp.cc:21-34 -- 64-bit XOR with 0x80.., n>>3 iterations (n multiple of 8) */
void orc_convtosigned(const uint8_t *in, uint8_t *out, int n)
{
    int N = n >> 3;
    for (int i = 0; i < N; i++) {
        uint64_t v;
        memcpy(&v, in + 8 * (size_t)i, 8);
        v ^= 0x8080808080808080ULL;
        memcpy(out + 8 * (size_t)i, &v, 8);
    }
}

/* src/cdsp.cc:36-44 -> volk_8i_s32f_convert_32f(out,in,127.0f,n), generic kernel:
 * iScalar = 1.0/scalar (rounded to float); out = (float)in * iScalar                        */
void orc_convtofloat(float *out, const int8_t *s8, int n)
{
    const float iScalar = (float)(1.0 / 127.0f);
    for (int i = 0; i < n; i++) out[i] = (float)s8[i] * iScalar;
}

/* src/cdsp.cc:51-54 -> volk_32f_s32f_convert_8i(out,in,127.0f,2n), generic kernel:
 * r = in*scalar; clamp to [-128,127]; (int8)rintf(r) (round-half-even)                      */
void orc_convto8bit(int8_t *out, const float *in, int nfloats)
{
    for (int i = 0; i < nfloats; i++) {
        float r = in[i] * 127.0f;
        if (r > 127.0f) r = 127.0f;
        else if (r < -128.0f) r = -128.0f;
        if (r != r) r = 0.0f;           /* NaN: reference is UB; both sides emit 0 */
        out[i] = (int8_t)rintf(r);
    }
}

/* src/cdsp.cc:46-49 -> volk_32fc_s32fc_multiply_32fc.  Rounding follows volk's x86 AVX+FMA
 * kernel (the one volk dispatches to on the hosts this runs on): tmp = swap(a)*im(s);
 * out = fmaddsub(a, re(s), tmp).                                                            */
void orc_scalarmul(orc_cf32 *out, const orc_cf32 *in, orc_cf32 s, int n)
{
    for (int i = 0; i < n; i++) {
        float ar = in[i].re, ai = in[i].im;
        float t0 = ai * s.im, t1 = ar * s.im;
        out[i].re = fmaf(ar, s.re, -t0);
        out[i].im = fmaf(ai, s.re, t1);
    }
}

/* src/cdsp.cc:61-66, src/csdrdevice.cc:62 -> volk_32fc_x2_conjugate_dot_prod_32fc generic:
 * two interleaved float accumulators over pairs of points, summed at the end.               */
orc_cf32 orc_conj_dotproduct(const orc_cf32 *a, const orc_cf32 *b, int n)
{
    const float *in = (const float *)a, *tp = (const float *)b;
    float s0r = 0, s0i = 0, s1r = 0, s1i = 0;
    int pairs = n >> 1;
    for (int i = 0; i < pairs; i++) {
        float p0 = in[0] * tp[0], p1 = in[1] * tp[1];
        float p2 = -in[0] * tp[1], p3 = in[1] * tp[0];
        float p4 = in[2] * tp[2], p5 = in[3] * tp[3];
        float p6 = -in[2] * tp[3], p7 = in[3] * tp[2];
        float q0 = p0 + p1, q1 = p2 + p3, q2 = p4 + p5, q3 = p6 + p7;
        s0r += q0; s0i += q1; s1r += q2; s1i += q3;
        in += 4; tp += 4;
    }
    orc_cf32 res; res.re = s0r + s1r; res.im = s0i + s1i;
    if (n & 1) {
        float xr = a[n-1].re, xi = a[n-1].im, yr = b[n-1].re, yi = b[n-1].im;
        float t0 = xr * yr, t1 = xi * yi, t2 = xi * yr, t3 = xr * yi;
        float u0 = t0 + t1, u1 = t2 - t3;
        res.re += u0; res.im += u1;
    }
    return res;
}

/* src/cdsp.cc:68-73 -> volk_32f_x2_dot_prod_32f generic (single accumulator) */
float orc_rms_f(const float *in, int n)
{
    float acc = 0;
    for (int i = 0; i < n; i++) { float p = in[i] * in[i]; acc += p; }
    return sqrtf(acc / n);
}

/* src/cdsp.cc:75-78 */
float orc_rms_c(const orc_cf32 *in, int n)
{
    orc_cf32 r = orc_conj_dotproduct(in, in, n);
    return sqrtf(r.re / n);
}

/* src/cdsp.cc:90-98 */
float orc_crestfactor(const float *in, int n)
{
    uint32_t idx = orc_indexofmax(in, n);
    return in[idx] / orc_rms_f(in, n);
}

/* src/cdsp.cc:100-103 -> volk_32fc_magnitude_squared_32f generic: re*re + im*im */
void orc_magsquared(float *out, const orc_cf32 *in, int n)
{
    for (int i = 0; i < n; i++) {
        float a = in[i].re * in[i].re, b = in[i].im * in[i].im;
        out[i] = a + b;
    }
}

/* src/cdsp.cc:105-108 -> volk_32fc_x2_multiply_conjugate_32fc generic: a * conj(b) */
void orc_conjugatemul(orc_cf32 *out, const orc_cf32 *a, const orc_cf32 *b, int n)
{
    for (int i = 0; i < n; i++) {
        float ar = a[i].re, ai = a[i].im, br = b[i].re, bi = b[i].im;
        float t0 = ar * br, t1 = ai * bi, t2 = ai * br, t3 = ar * bi;
        out[i].re = t0 + t1;
        out[i].im = t2 - t3;
    }
}

/* src/cdsp.cc:135-139 -> volk_32f_index_max_32u generic: first index of the strict maximum */
uint32_t orc_indexofmax(const float *in, int n)
{
    uint32_t idx = 0; float m = in[0];
    for (int i = 1; i < n; i++) if (in[i] > m) { m = in[i]; idx = (uint32_t)i; }
    return idx;
}
