/*
 * cdsp_oracle_fft.c -- float32 c2c DFT for the oracle (TEST INFRASTRUCTURE ONLY).
 *
 * Stands in for fftw3f (un-vendored, unpinned; call sites src/ccoherent.cc:87-93,174,182 and
 * src/cdsp.cc:110-120): forward sign -1, backward sign +1, both unnormalised, out-of-place.
 * Algorithm: Stockham autosort, radix-4 passes plus one radix-2 pass when log2(n) is odd; AVX2 + FMA
 * kernels for every pass of a transform of 16 points or more.
 * Twiddles are computed in double and rounded once to float.
 */
#include "cdsp_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

#define ORC_MAX_PLANS 64
typedef struct { int n; orc_cf32 *tw; orc_cf32 *tw3; } orc_plan;   /* tw3: w^p, w^2p, w^3p (p < n/4) as three contiguous arrays */
static orc_plan g_plans[ORC_MAX_PLANS];
static int g_nplans = 0;
static pthread_mutex_t g_plan_mtx = PTHREAD_MUTEX_INITIALIZER;

static const orc_plan *get_plan(int n)
{
    const orc_plan *pl = NULL;
    pthread_mutex_lock(&g_plan_mtx);
    for (int i = 0; i < g_nplans; i++) if (g_plans[i].n == n) pl = &g_plans[i];
    if (!pl && g_nplans < ORC_MAX_PLANS) {
        orc_cf32 *t = (orc_cf32 *)malloc(sizeof(orc_cf32) * (size_t)n);
        for (int k = 0; k < n; k++) {
            double a = -2.0 * M_PI * (double)k / (double)n;
            t[k].re = (float)cos(a); t[k].im = (float)sin(a);
        }
        const int n1 = n >> 2;
        orc_cf32 *t3 = (orc_cf32 *)malloc(sizeof(orc_cf32) * (size_t)(3 * n1 + 4));
        for (int j = 1; j <= 3; j++)
            for (int p = 0; p < n1; p++) t3[(j - 1) * n1 + p] = t[(j * p) % n];
        g_plans[g_nplans].n = n; g_plans[g_nplans].tw = t; g_plans[g_nplans].tw3 = t3;
        pl = &g_plans[g_nplans]; g_nplans++;
    }
    pthread_mutex_unlock(&g_plan_mtx);
    return pl;
}

/* ---- AVX2 + FMA kernels (x86-64-v3, what the Makefile targets) ---------------------------------
 * A complex float is one 64-bit lane; a __m256 holds four of them {re0 im0 re1 im1 ...}.
 * round 2: the scalar passes left the s = 1 / 4 / 16 inner loops unvectorised (224 us per 16384-point
 * transform against 111 us for scipy's pocketfft); these run at pocketfft speed or better. */
#ifndef __AVX2__
/* There is deliberately no scalar variant: the committed goldens (the npz files under tests/golden) carry THIS code's
 * rounding (FMA contraction inside the intrinsics), and a second code path with different rounding would
 * be a second, unpinned checker.  oracle/Makefile builds with -march=x86-64-v3. */
#error "cdsp_oracle_fft.c needs AVX2 + FMA (build with -march=x86-64-v3 as oracle/Makefile does)"
#endif
#include <immintrin.h>

static inline __m256 cmul4(__m256 a, __m256 wr, __m256 wi)       /* a * (wr + j wi), lane-wise */
{
    __m256 asw = _mm256_permute_ps(a, 0xB1);                    /* {im, re} */
    return _mm256_fmaddsub_ps(a, wr, _mm256_mul_ps(asw, wi));  /* re: a.re wr - a.im wi ; im: a.im wr + a.re wi */
}
static inline __m256 mulj4(__m256 a)                             /* j * a = (-im, re) */
{
    const __m256 sgn = _mm256_castsi256_ps(_mm256_set_epi32(0, (int)0x80000000u, 0, (int)0x80000000u, 0, (int)0x80000000u, 0, (int)0x80000000u));
    return _mm256_xor_ps(_mm256_permute_ps(a, 0xB1), sgn);
}
#define BFLY4(a, b, c, d, sign, o0, t1, t2, t3) do {                                   \
        __m256 apc = _mm256_add_ps(a, c), amc = _mm256_sub_ps(a, c);                    \
        __m256 bpd = _mm256_add_ps(b, d), jb = mulj4(_mm256_sub_ps(b, d));              \
        o0 = _mm256_add_ps(apc, bpd); t2 = _mm256_sub_ps(apc, bpd);                     \
        if (sign < 0) { t1 = _mm256_sub_ps(amc, jb); t3 = _mm256_add_ps(amc, jb); }     \
        else          { t1 = _mm256_add_ps(amc, jb); t3 = _mm256_sub_ps(amc, jb); }     \
    } while (0)

/* radix-4 Stockham pass with stride s >= 4 (a multiple of 4): vectors run along q */
static void pass4_q(const orc_cf32 *x, orc_cf32 *y, int n, int nc, int s, const orc_cf32 *tw, int sign)
{
    const int n1 = nc >> 2, tstep = n / nc;
    const float cs = sign > 0 ? -1.0f : 1.0f;
    for (int p = 0; p < n1; p++) {
        const orc_cf32 w1 = tw[p * tstep], w2 = tw[2 * p * tstep], w3 = tw[3 * p * tstep];
        const __m256 w1r = _mm256_set1_ps(w1.re), w1i = _mm256_set1_ps(cs * w1.im);
        const __m256 w2r = _mm256_set1_ps(w2.re), w2i = _mm256_set1_ps(cs * w2.im);
        const __m256 w3r = _mm256_set1_ps(w3.re), w3i = _mm256_set1_ps(cs * w3.im);
        const float *xa = (const float *)(x + (size_t)s * p), *xb = xa + 2 * (size_t)s * n1,
                    *xc = xb + 2 * (size_t)s * n1, *xd = xc + 2 * (size_t)s * n1;
        float *y0 = (float *)(y + (size_t)s * 4 * p), *y1 = y0 + 2 * s, *y2 = y1 + 2 * s, *y3 = y2 + 2 * s;
        for (int q = 0; q < 2 * s; q += 8) {
            __m256 a = _mm256_loadu_ps(xa + q), b = _mm256_loadu_ps(xb + q), c = _mm256_loadu_ps(xc + q), d = _mm256_loadu_ps(xd + q);
            __m256 o0, t1, t2, t3;
            BFLY4(a, b, c, d, sign, o0, t1, t2, t3);
            _mm256_storeu_ps(y0 + q, o0);
            if (p == 0) { _mm256_storeu_ps(y1 + q, t1); _mm256_storeu_ps(y2 + q, t2); _mm256_storeu_ps(y3 + q, t3); }
            else {
                _mm256_storeu_ps(y1 + q, cmul4(t1, w1r, w1i));
                _mm256_storeu_ps(y2 + q, cmul4(t2, w2r, w2i));
                _mm256_storeu_ps(y3 + q, cmul4(t3, w3r, w3i));
            }
        }
    }
}

/* first radix-4 pass (s = 1, nc = n): vectors run along p; twk = per-p twiddles {w1[p], w2[p], w3[p]} laid out
 * as three contiguous arrays; the four outputs of a butterfly are adjacent in y: 4 x 4 transpose in registers */
static void pass4_first(const orc_cf32 *x, orc_cf32 *y, int n, const orc_cf32 *tw3, int sign)
{
    const int n1 = n >> 2;
    const float *xa = (const float *)x, *xb = xa + 2 * (size_t)n1, *xc = xb + 2 * (size_t)n1, *xd = xc + 2 * (size_t)n1;
    const float *w1 = (const float *)tw3, *w2 = w1 + 2 * (size_t)n1, *w3 = w2 + 2 * (size_t)n1;
    const __m256 cs = sign > 0 ? _mm256_castsi256_ps(_mm256_set_epi32((int)0x80000000u, 0, (int)0x80000000u, 0, (int)0x80000000u, 0, (int)0x80000000u, 0))
                               : _mm256_setzero_ps();           /* conjugates the twiddles for the backward transform */
    float *yo = (float *)y;
    for (int p = 0; p < n1; p += 4) {
        __m256 a = _mm256_loadu_ps(xa + 2 * p), b = _mm256_loadu_ps(xb + 2 * p), c = _mm256_loadu_ps(xc + 2 * p), d = _mm256_loadu_ps(xd + 2 * p);
        __m256 o0, t1, t2, t3;
        BFLY4(a, b, c, d, sign, o0, t1, t2, t3);
        __m256 v1 = _mm256_xor_ps(_mm256_loadu_ps(w1 + 2 * p), cs), v2 = _mm256_xor_ps(_mm256_loadu_ps(w2 + 2 * p), cs), v3 = _mm256_xor_ps(_mm256_loadu_ps(w3 + 2 * p), cs);
        t1 = cmul4(t1, _mm256_moveldup_ps(v1), _mm256_movehdup_ps(v1));
        t2 = cmul4(t2, _mm256_moveldup_ps(v2), _mm256_movehdup_ps(v2));
        t3 = cmul4(t3, _mm256_moveldup_ps(v3), _mm256_movehdup_ps(v3));
        /* rows = outputs j (o0,t1,t2,t3), columns = p..p+3  ->  y[4 p' + j] */
        __m256d r0 = _mm256_castps_pd(o0), r1 = _mm256_castps_pd(t1), r2 = _mm256_castps_pd(t2), r3 = _mm256_castps_pd(t3);
        __m256d l01 = _mm256_unpacklo_pd(r0, r1), h01 = _mm256_unpackhi_pd(r0, r1);   /* {j0p0 j1p0 j0p2 j1p2}, {j0p1 j1p1 j0p3 j1p3} */
        __m256d l23 = _mm256_unpacklo_pd(r2, r3), h23 = _mm256_unpackhi_pd(r2, r3);
        _mm256_storeu_pd((double *)(yo + 8 * p),      _mm256_permute2f128_pd(l01, l23, 0x20));   /* p+0 */
        _mm256_storeu_pd((double *)(yo + 8 * p + 8),  _mm256_permute2f128_pd(h01, h23, 0x20));   /* p+1 */
        _mm256_storeu_pd((double *)(yo + 8 * p + 16), _mm256_permute2f128_pd(l01, l23, 0x31));   /* p+2 */
        _mm256_storeu_pd((double *)(yo + 8 * p + 24), _mm256_permute2f128_pd(h01, h23, 0x31));   /* p+3 */
    }
}

/* scalar radix-4 pass: small transforms (n < 16) and the s = 1 / 2 passes the vector kernels do not cover */
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
            orc_cf32 jbmd = { -bmd.im, bmd.re };
            orc_cf32 t1, t3;
            if (sign < 0) { t1.re = amc.re - jbmd.re; t1.im = amc.im - jbmd.im;
                            t3.re = amc.re + jbmd.re; t3.im = amc.im + jbmd.im; }
            else          { t1.re = amc.re + jbmd.re; t1.im = amc.im + jbmd.im;
                            t3.re = amc.re - jbmd.re; t3.im = amc.im - jbmd.im; }
            orc_cf32 t2 = { apc.re - bpd.re, apc.im - bpd.im };
            y0[q].re = apc.re + bpd.re; y0[q].im = apc.im + bpd.im;
            /* same rounding as cmul4: im/re = fma(x, w.re, -+(x' * w.im)) */
            y1[q].re = fmaf(t1.re, w1.re, -(t1.im * w1.im)); y1[q].im = fmaf(t1.im, w1.re, t1.re * w1.im);
            y2[q].re = fmaf(t2.re, w2.re, -(t2.im * w2.im)); y2[q].im = fmaf(t2.im, w2.re, t2.re * w2.im);
            y3[q].re = fmaf(t3.re, w3.re, -(t3.im * w3.im)); y3[q].im = fmaf(t3.im, w3.re, t3.re * w3.im);
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
    const orc_plan *pl = get_plan(n);
    const orc_cf32 *tw = pl->tw;
    /* per-thread scratch, grown on demand (no malloc per call) */
    static __thread orc_cf32 *tmp = NULL; static __thread int tmp_n = 0;
    if (tmp_n < n) {
        /* aligned_alloc wants a size that is a multiple of the alignment (n < 8 would not be).  The buffer
         * lives as long as its thread (a few worker threads of the harness; never per call). */
        const size_t bytes = (sizeof(orc_cf32) * (size_t)n + 63) / 64 * 64;
        free(tmp); tmp = (orc_cf32 *)aligned_alloc(64, bytes); tmp_n = n;
    }
    int passes = 0;
    for (int nc = n; nc > 1; nc = (nc >= 4) ? nc >> 2 : nc >> 1) passes++;
    /* ping-pong so that the last pass lands in out; the first pass reads in (never written) */
    orc_cf32 *bufs[2] = { out, tmp };
    int dst = (passes & 1) ? 0 : 1;
    const orc_cf32 *src = in;
    int s = 1;
    for (int nc = n; nc > 1; ) {
        orc_cf32 *y = bufs[dst];
        if (nc >= 4) {
            if (s == 1 && n >= 16) pass4_first(src, y, n, pl->tw3, sign);
            else if (s >= 4) pass4_q(src, y, n, nc, s, tw, sign);
            else pass4(src, y, n, nc, s, tw, sign);
            nc >>= 2; s <<= 2;
        } else { pass2(src, y, s); nc >>= 1; s <<= 1; }
        src = y; dst ^= 1;
    }
}
