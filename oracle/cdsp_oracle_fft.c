/*
 * cdsp_oracle_fft.c -- float32 c2c DFT for the oracle (TEST INFRASTRUCTURE ONLY).
 *
 * Stands in for fftw3f (un-vendored, unpinned; call sites src/ccoherent.cc:87-93,174,182 and
 * src/cdsp.cc:110-120): forward sign -1, backward sign +1, both unnormalised, out-of-place.
 * Algorithm: Stockham autosort, radix-4 passes plus one radix-2 pass when log2(n) is odd.
 * Twiddles are computed in double and rounded once to float.
 */
#include "cdsp_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

#define ORC_MAX_PLANS 64
typedef struct { int n; orc_cf32 *tw; } orc_plan;
static orc_plan g_plans[ORC_MAX_PLANS];
static int g_nplans = 0;
static pthread_mutex_t g_plan_mtx = PTHREAD_MUTEX_INITIALIZER;

static const orc_cf32 *get_twiddles(int n)
{
    const orc_cf32 *tw = NULL;
    pthread_mutex_lock(&g_plan_mtx);
    for (int i = 0; i < g_nplans; i++) if (g_plans[i].n == n) tw = g_plans[i].tw;
    if (!tw && g_nplans < ORC_MAX_PLANS) {
        orc_cf32 *t = (orc_cf32 *)malloc(sizeof(orc_cf32) * (size_t)n);
        for (int k = 0; k < n; k++) {
            double a = -2.0 * M_PI * (double)k / (double)n;
            t[k].re = (float)cos(a); t[k].im = (float)sin(a);
        }
        g_plans[g_nplans].n = n; g_plans[g_nplans].tw = t; g_nplans++;
        tw = t;
    }
    pthread_mutex_unlock(&g_plan_mtx);
    return tw;
}

static inline orc_cf32 cmul(orc_cf32 a, orc_cf32 w)
{
    orc_cf32 r;
    float t0 = a.re * w.re, t1 = a.im * w.im, t2 = a.re * w.im, t3 = a.im * w.re;
    r.re = t0 - t1; r.im = t2 + t3;
    return r;
}

/* one radix-4 Stockham pass: sub-transform length nc, stride s; x -> y */
static void pass4(const orc_cf32 *x, orc_cf32 *y, int n, int nc, int s, const orc_cf32 *tw, int sign)
{
    const int n1 = nc >> 2, tstep = n / nc;
    for (int p = 0; p < n1; p++) {
        orc_cf32 w1 = tw[p * tstep], w2 = tw[2 * p * tstep], w3 = tw[3 * p * tstep];
        if (sign > 0) { w1.im = -w1.im; w2.im = -w2.im; w3.im = -w3.im; }
        const orc_cf32 *xa = x + (size_t)s * p, *xb = xa + (size_t)s * n1,
                       *xc = xb + (size_t)s * n1, *xd = xc + (size_t)s * n1;
        orc_cf32 *y0 = y + (size_t)s * 4 * p, *y1 = y0 + s, *y2 = y1 + s, *y3 = y2 + s;
        for (int q = 0; q < s; q++) {
            orc_cf32 a = xa[q], b = xb[q], c = xc[q], d = xd[q];
            orc_cf32 apc = { a.re + c.re, a.im + c.im }, amc = { a.re - c.re, a.im - c.im };
            orc_cf32 bpd = { b.re + d.re, b.im + d.im }, bmd = { b.re - d.re, b.im - d.im };
            /* j*(b-d) */
            orc_cf32 jbmd = { -bmd.im, bmd.re };
            orc_cf32 t1, t3;
            if (sign < 0) { t1.re = amc.re - jbmd.re; t1.im = amc.im - jbmd.im;
                            t3.re = amc.re + jbmd.re; t3.im = amc.im + jbmd.im; }
            else          { t1.re = amc.re + jbmd.re; t1.im = amc.im + jbmd.im;
                            t3.re = amc.re - jbmd.re; t3.im = amc.im - jbmd.im; }
            orc_cf32 t2 = { apc.re - bpd.re, apc.im - bpd.im };
            y0[q].re = apc.re + bpd.re; y0[q].im = apc.im + bpd.im;
            y1[q] = cmul(t1, w1);
            y2[q] = cmul(t2, w2);
            y3[q] = cmul(t3, w3);
        }
    }
}

/* final radix-2 pass (nc == 2): no twiddles */
static void pass2(const orc_cf32 *x, orc_cf32 *y, int s)
{
    for (int q = 0; q < s; q++) {
        orc_cf32 a = x[q], b = x[q + s];
        y[q].re = a.re + b.re; y[q].im = a.im + b.im;
        y[q + s].re = a.re - b.re; y[q + s].im = a.im - b.im;
    }
}

void orc_fft(orc_cf32 *out, const orc_cf32 *in, int n, int sign)
{
    if (n == 1) { out[0] = in[0]; return; }
    const orc_cf32 *tw = get_twiddles(n);
    orc_cf32 *tmp = (orc_cf32 *)malloc(sizeof(orc_cf32) * (size_t)n);
    int passes = 0;
    for (int nc = n; nc > 1; nc = (nc >= 4) ? nc >> 2 : nc >> 1) passes++;
    /* ping-pong so that the last pass lands in out; the first pass reads in (never written) */
    orc_cf32 *bufs[2] = { out, tmp };
    int dst = (passes & 1) ? 0 : 1;
    const orc_cf32 *src = in;
    int s = 1;
    for (int nc = n; nc > 1; ) {
        orc_cf32 *y = bufs[dst];
        if (nc >= 4) { pass4(src, y, n, nc, s, tw, sign); nc >>= 2; s <<= 2; }
        else         { pass2(src, y, s); nc >>= 1; s <<= 1; }
        src = y; dst ^= 1;
    }
    free(tmp);
}
