/*
 * cdsp_oracle.h -- CPU restatement of the coherent-rtlsdr block sync + correction path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may call it,
 * and only as the checker / the timed CPU baseline.  The shipped path is libcoherent_cuda.so.
 *
 * Parity status: pinned against the reference's own translation units compiled from
 * /root/reference (oracle/Makefile -> oracle/_ref/libref_cdsp.so, see oracle/ref_harness.cc)
 * and against tests/golden/ vectors produced by that build.  The reference itself ships no
 * tests or golden vectors.
 *
 * Third-party arithmetic restated here (absent from /root/reference, versions unpinned by the
 * reference -- install_on_rpi clones master): volk generic kernels and fftw3f c2c transforms.
 *
 * All citations are file:line in /root/reference.
 */
#ifndef CDSP_ORACLE_H
#define CDSP_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float re, im; } orc_cf32;

/* --- cdsp primitives (src/cdsp.cc) ------------------------------------------------------ */
void  orc_convtosigned(const uint8_t *in, uint8_t *out, int n);               /* cdsp.cc:21-34  */
void  orc_convtofloat(float *out, const int8_t *s8, int n);                   /* cdsp.cc:36-44  */
void  orc_convto8bit(int8_t *out, const float *in, int nfloats);              /* cdsp.cc:51-54  */
void  orc_scalarmul(orc_cf32 *out, const orc_cf32 *in, orc_cf32 s, int n);    /* cdsp.cc:46-49  */
orc_cf32 orc_conj_dotproduct(const orc_cf32 *a, const orc_cf32 *b, int n);    /* cdsp.cc:61-66  */
float orc_rms_f(const float *in, int n);                                      /* cdsp.cc:68-73  */
float orc_rms_c(const orc_cf32 *in, int n);                                   /* cdsp.cc:75-78  */
float orc_crestfactor(const float *in, int n);                                /* cdsp.cc:90-98  */
void  orc_magsquared(float *out, const orc_cf32 *in, int n);                  /* cdsp.cc:100-103 */
void  orc_conjugatemul(orc_cf32 *out, const orc_cf32 *a, const orc_cf32 *b, int n); /* :105-108 */
uint32_t orc_indexofmax(const float *in, int n);                              /* cdsp.cc:135-139 */
/* unnormalised c2c DFT, sign=-1 forward / +1 backward (fftw semantics, ccoherent.cc:87-93) */
void  orc_fft(orc_cf32 *out, const orc_cf32 *in, int n, int sign);

/* --- per-block path (src/ccoherent.cc:245-294 schedule) --------------------------------- */
typedef struct {
    uint32_t n_signal;       /* C: signal channels (ref excluded)                           */
    uint32_t blocksize;      /* B: bytes per channel-block = 2L (src/main.cc:167)           */
    float    alpha;          /* IIR weight, reference hard-codes 0.5 (csdrdevice.cc:60)     */
    int      noheader;       /* -R raw mode (cpacketizer.cc:91-96)                          */
    int      apply_shift;    /* 0 = reference behaviour (no software shift)      */
    int      int_dot;        /* 0 = volk float dot (reference); 1 = exact integer dot spec  */
    int      max_lag_slots;  /* reference nfft-1 = 7 (ccoherent.cc:124); <=0 = unlimited    */
} orc_config;

typedef struct {
    orc_cf32 phasecorr;      /* csdrdevice.h:81 */
    orc_cf32 phasecorrprev;  /* csdrdevice.h:82 */
    float    lag;            /* lagpoint.lag, csdrdevice.h:42-54 */
    float    mag;
    int32_t  held_lag;       /* software-shift extension: integer lag applied while tracking */
} orc_chan_state;

size_t orc_packet_len(const orc_config *cfg);     /* bytes actually written: hdr + columns    */
void   orc_state_init(orc_chan_state *st, uint32_t n);   /* csdrdevice.cc:36-40                */

/*
 * One block through the threadf schedule.  chan[0] = ref, chan[1..C] = signal channels, raw
 * offset-binary u8 as delivered by librtlsdr (input_is_signed=1 if the a-1 XOR already ran).
 * lag_request: C+1 flags (entry 0 ignored).  st: C+1 states (entry 0 unused except lag/mag=0).
 * packet: orc_packet_len() bytes; hdr0.globalseqn := globalseqn.
 * cf32_out: optional (C+1) x L corrected float matrix (column 0 = converted ref), or NULL.
 */
void orc_process_block(const orc_config *cfg, orc_chan_state *st,
                       const uint8_t *const *chan, const uint32_t *readcnt,
                       const uint8_t *lag_request, int refnoise_enabled, int input_is_signed,
                       uint32_t globalseqn, int8_t *packet, orc_cf32 *cf32_out);

/* contiguous batch: iq = [n_blocks][C+1][B]; flags [n_blocks][C+1] or NULL (=all 1 / all 0 by
 * lag_mode: 1 SYNC, 0 TRACK); refnoise [n_blocks] or NULL (= enabled). Outputs per block:
 * packets [n_blocks][packet_len], lag/mag [n_blocks][C+1], phase [n_blocks][C+1] (cf32). */
void orc_process_batch(const orc_config *cfg, orc_chan_state *st, const uint8_t *iq,
                       uint32_t n_blocks, const uint8_t *lag_request, int lag_mode,
                       const uint8_t *refnoise, int input_is_signed, uint32_t readcnt0,
                       int8_t *packets, float *lag, float *mag, orc_cf32 *phase,
                       orc_cf32 *cf32_out);

/* correlation only (ccoherent.cc:154-239) for one signal/ref pair of converted floats */
void orc_computelag_pair(const orc_cf32 *x /*L*/, const orc_cf32 *r /*L*/, uint32_t L,
                         uint32_t *idx, float *lag, float *mag, float *peak_margin);

#ifdef __cplusplus
}
#endif
#endif
