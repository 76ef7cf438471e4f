/*
 * cdsp_oracle_path.c -- the per-block schedule of ccoherent::threadf restated on the CPU.
 * TEST INFRASTRUCTURE ONLY (see cdsp_oracle.h).  Citations are file:line in /root/reference.
 */
#include "cdsp_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

size_t orc_packet_len(const orc_config *cfg)
{
    /* bytes actually written per packet (cpacketizer.cc:95 allocates twice that) */
    size_t nch = (size_t)cfg->n_signal + 1;
    size_t hdr = cfg->noheader ? 0 : 16 + 4 * nch;      /* struct hdr0 + readcnt[nch]          */
    return hdr + nch * (size_t)cfg->blocksize;
}

/* csdrdevice.cc:36-40 */
void orc_state_init(orc_chan_state *st, uint32_t n)
{
    for (uint32_t i = 0; i < n; i++) {
        st[i].phasecorr.re = 1.0f; st[i].phasecorr.im = 0.0f;
        st[i].phasecorrprev = st[i].phasecorr;
        st[i].lag = 0.0f; st[i].mag = 0.0f; st[i].held_lag = 0;
    }
}

/*
 * ccoherent::computelag for one queued channel (ccoherent.cc:154-239), using the buffer
 * layout of crtlsdr::convtofloat / crefsdr::convtofloat (crtlsdr.cc:205-218): the signal
 * occupies [0,L) and the reference [L,2L) of a zero-padded N=2L buffer.
 */
void orc_computelag_pair(const orc_cf32 *x, const orc_cf32 *r, uint32_t L,
                         uint32_t *idx_out, float *lag_out, float *mag_out, float *margin_out)
{
    const uint32_t N = 2 * L;
    orc_cf32 *sx = (orc_cf32 *)calloc(N, sizeof(orc_cf32));
    orc_cf32 *sr = (orc_cf32 *)calloc(N, sizeof(orc_cf32));
    orc_cf32 *fx = (orc_cf32 *)malloc(N * sizeof(orc_cf32));
    orc_cf32 *fr = (orc_cf32 *)malloc(N * sizeof(orc_cf32));
    float *msq = (float *)malloc(N * sizeof(float));
    memcpy(sx, x, L * sizeof(orc_cf32));              /* signal -> first half  */
    memcpy(sr + L, r, L * sizeof(orc_cf32));          /* ref    -> second half */
    orc_fft(fx, sx, (int)N, -1);                      /* ccoherent.cc:174 */
    orc_fft(fr, sr, (int)N, -1);
    orc_conjugatemul(sx, fx, fr, (int)N);             /* ccoherent.cc:177-179 */
    orc_fft(fx, sx, (int)N, +1);                      /* ccoherent.cc:182 (unnormalised) */
    orc_magsquared(msq, fx, (int)N);                  /* ccoherent.cc:185 */
    uint32_t idx = orc_indexofmax(msq, (int)N);       /* ccoherent.cc:192 */
    float mag = sqrtf(msq[idx] / (float)L);           /* ccoherent.cc:203-204 */
    float lag = (float)idx;                           /* ccoherent.cc:218/221 (fractional part disabled) */
    lag -= (float)L;                                  /* ccoherent.cc:232 */
    if (margin_out) {
        float second = 0.0f;
        for (uint32_t i = 0; i < N; i++) if (i != idx && msq[i] > second) second = msq[i];
        *margin_out = msq[idx] > 0.0f ? second / msq[idx] : 1.0f;
    }
    *idx_out = idx; *lag_out = lag; *mag_out = mag;
    free(sx); free(sr); free(fx); free(fr); free(msq);
}

/* csdrdevice::est_phasecorrect (csdrdevice.cc:58-69), float dot as the reference */
static orc_cf32 est_phase_float(const orc_config *cfg, orc_chan_state *st,
                                const orc_cf32 *x, const orc_cf32 *r, uint32_t L)
{
    orc_cf32 corr = orc_conj_dotproduct(x, r, (int)L);
    float inv = 1.0f / hypotf(corr.re, corr.im);      /* std::abs(complex<float>) */
    orc_cf32 p = { corr.re * inv, -corr.im * inv };
    float a = cfg->alpha, b = 1.0f - cfg->alpha;
    float t0 = a * p.re, t1 = b * st->phasecorrprev.re;
    float t2 = a * p.im, t3 = b * st->phasecorrprev.im;
    p.re = t0 + t1; p.im = t2 + t3;
    st->phasecorr = p; st->phasecorrprev = p;
    return p;
}

/*
 * Integer-dot variant (the GPU path's specification, DESIGN.md "phase estimate"): the samples
 * are int8-valued, so sum x*conj(r) is formed exactly in integers on the raw bytes and only
 * the direction conj(S)/|S| is taken in float with individually rounded operations.
 */
static orc_cf32 est_phase_int(const orc_config *cfg, orc_chan_state *st,
                              const int8_t *xs /*shifted, 2L*/, const int8_t *rs, uint32_t L)
{
    int64_t sre = 0, sim = 0;
    for (uint32_t n = 0; n < L; n++) {
        int32_t xr = xs[2*n], xi = xs[2*n+1], rr = rs[2*n], ri = rs[2*n+1];
        sre += xr * rr + xi * ri;
        sim += xi * rr - xr * ri;
    }
    float cr = (float)sre, ci = (float)sim;
    float t = ci * ci;
    float m2 = fmaf(cr, cr, t);
    float m = sqrtf(m2);
    float inv = 1.0f / m;
    orc_cf32 p = { cr * inv, -(ci * inv) };
    float a = cfg->alpha, b = 1.0f - cfg->alpha;
    float t0 = a * p.re, t1 = b * st->phasecorrprev.re;
    float t2 = a * p.im, t3 = b * st->phasecorrprev.im;
    p.re = t0 + t1; p.im = t2 + t3;
    st->phasecorr = p; st->phasecorrprev = p;
    return p;
}

void orc_process_block(const orc_config *cfg, orc_chan_state *st,
                       const uint8_t *const *chan, const uint32_t *readcnt,
                       const uint8_t *lag_request, int refnoise_enabled, int input_is_signed,
                       uint32_t globalseqn, int8_t *packet, orc_cf32 *cf32_out)
{
    const uint32_t B = cfg->blocksize, L = B >> 1, C = cfg->n_signal, nch = C + 1;
    const size_t hdr = cfg->noheader ? 0 : 16 + 4 * (size_t)nch;
    int8_t *rs = (int8_t *)malloc(B), *xs = (int8_t *)malloc(B), *xsh = (int8_t *)malloc(B);
    orc_cf32 *rf = (orc_cf32 *)malloc(L * sizeof(orc_cf32));
    orc_cf32 *xf = (orc_cf32 *)malloc(L * sizeof(orc_cf32));
    int8_t *scratch = NULL;
    if (!packet) packet = scratch = (int8_t *)malloc(orc_packet_len(cfg));

    if (!cfg->noheader) {                             /* cpacketizer.cc:110-118, struct hdr0 */
        uint32_t h[4] = { globalseqn, nch, L, 0 };
        memcpy(packet, h, 16);
    }
    /* reference channel: ccoherent.cc:251-254 */
    if (input_is_signed) memcpy(rs, chan[0], B);
    else orc_convtosigned(chan[0], (uint8_t *)rs, (int)B);          /* common.h:121 */
    orc_convtofloat((float *)rf, rs, (int)B);                        /* crtlsdr.cc:215-218 */
    if (!cfg->noheader) memcpy(packet + 16, &readcnt[0], 4);         /* cpacketizer.cc:142 */
    memcpy(packet + hdr, rs, B);                                     /* cpacketizer.cc:151 */
    if (cf32_out) memcpy(cf32_out, rf, L * sizeof(orc_cf32));

    int slots_used = 0;
    for (uint32_t c = 1; c <= C; c++) {                              /* ccoherent.cc:262-283 */
        if (input_is_signed) memcpy(xs, chan[c], B);
        else orc_convtosigned(chan[c], (uint8_t *)xs, (int)B);
        orc_convtofloat((float *)xf, xs, (int)B);                    /* crtlsdr.cc:205-207 */

        /* lag is measured on the un-shifted, un-rotated samples (ccoherent.cc:266-267) */
        int measure = lag_request && lag_request[c] &&
                      (cfg->max_lag_slots <= 0 || slots_used < cfg->max_lag_slots);
        if (measure) {
            uint32_t idx; float lag, mag;
            orc_computelag_pair(xf, rf, L, &idx, &lag, &mag, NULL);
            st[c].lag = lag; st[c].mag = mag;                        /* csdrdevice.h:138-151 */
            st[c].held_lag = (int32_t)idx - (int32_t)L;
            slots_used++;
        }
        /* software shift (not in the reference): x'[n] = x[n+lag] or 0 */
        int32_t sh = cfg->apply_shift ? st[c].held_lag : 0;
        if (sh != 0) {
            for (uint32_t n = 0; n < L; n++) {
                int64_t m = (int64_t)n + sh;
                if (m >= 0 && m < (int64_t)L) { xsh[2*n] = xs[2*m]; xsh[2*n+1] = xs[2*m+1]; }
                else { xsh[2*n] = 0; xsh[2*n+1] = 0; }
            }
            orc_convtofloat((float *)xf, xsh, (int)B);
        } else memcpy(xsh, xs, B);

        if (refnoise_enabled) {                                      /* ccoherent.cc:271-273 */
            if (cfg->int_dot) est_phase_int(cfg, &st[c], xsh, rs, L);
            else est_phase_float(cfg, &st[c], xf, rf, L);
        }
        orc_scalarmul(xf, xf, st[c].phasecorr, (int)L);              /* ccoherent.cc:275 */
        if (!cfg->noheader) memcpy(packet + 16 + 4 * (size_t)c, &readcnt[c], 4);
        orc_convto8bit(packet + hdr + (size_t)c * B, (const float *)xf, (int)B); /* cpacketizer.cc:158-172 */
        if (cf32_out) memcpy(cf32_out + (size_t)c * L, xf, L * sizeof(orc_cf32));
    }
    free(rs); free(xs); free(xsh); free(rf); free(xf); free(scratch);
}

void orc_process_batch(const orc_config *cfg, orc_chan_state *st, const uint8_t *iq,
                       uint32_t n_blocks, const uint8_t *lag_request, int lag_mode,
                       const uint8_t *refnoise, int input_is_signed, uint32_t readcnt0,
                       int8_t *packets, float *lag, float *mag, orc_cf32 *phase,
                       orc_cf32 *cf32_out)
{
    const uint32_t B = cfg->blocksize, L = B >> 1, nch = cfg->n_signal + 1;
    const size_t plen = orc_packet_len(cfg);
    const uint8_t **chan = (const uint8_t **)malloc(sizeof(uint8_t *) * nch);
    uint32_t *rc = (uint32_t *)malloc(4 * nch);
    uint8_t *flags = (uint8_t *)malloc(nch);
    for (uint32_t b = 0; b < n_blocks; b++) {
        for (uint32_t c = 0; c < nch; c++) {
            chan[c] = iq + ((size_t)b * nch + c) * B;
            rc[c] = readcnt0 + b;
            flags[c] = lag_request ? lag_request[(size_t)b * nch + c] : (uint8_t)(lag_mode != 0);
        }
        orc_process_block(cfg, st, chan, rc, flags, refnoise ? refnoise[b] : 1, input_is_signed,
                          b, packets ? packets + (size_t)b * plen : NULL,
                          cf32_out ? cf32_out + (size_t)b * nch * L : NULL);
        for (uint32_t c = 0; c < nch; c++) {
            if (lag)   lag[(size_t)b * nch + c] = st[c].lag;
            if (mag)   mag[(size_t)b * nch + c] = st[c].mag;
            if (phase) phase[(size_t)b * nch + c] = st[c].phasecorr;
        }
    }
    free(chan); free(rc); free(flags);
}
