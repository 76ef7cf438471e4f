/*
 * oracle/shim/shim_impl.cc -- implementations behind the declaration shims (volk, fftw3f,
 * librtlsdr).  TEST INFRASTRUCTURE ONLY: linked with the reference's own translation units into
 * oracle/_ref/libref_cdsp.so so that the reference's schedule (ccoherent::threadf), buffer
 * layout, lag/phase arithmetic and packet writer run hardware-free as the parity oracle and the
 * CPU baseline.  Nothing here is part of the product.
 *
 * VOLK kernels: the published generic-C semantics of each kernel the reference calls (VOLK is
 * unpinned by the reference, install_on_rpi clones master).  Where VOLK's x86 SIMD kernels
 * round differently from the generic one, the variant is named at the function.
 * FFTW: forward -1 / backward +1, unnormalised, executed by the oracle's own float32 Stockham
 * transform (cdsp_oracle_fft.c) -- "any correct float32 c2c FFT" (SURVEY 8c).
 * Build with -ffp-contract=off: every float expression rounds where it is written.
 */
#include <volk/volk.h>
#include <fftw3.h>
#include <rtl-sdr.h>

#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#include "../cdsp_oracle.h"   /* orc_fft */

/* ------------------------------------------------------------------------------- volk ---- */
size_t volk_get_alignment(void) { return 64; }
void *volk_malloc(size_t size, size_t alignment)
{
    void *p = nullptr;
    if (alignment < sizeof(void *)) alignment = sizeof(void *);
    if (posix_memalign(&p, alignment, size ? size : alignment) != 0) return nullptr;
    return p;
}
void volk_free(void *p) { free(p); }

/* volk_8i_s32f_convert_32f_generic: iScalar = 1.0 / scalar; out = (float)in * iScalar */
void volk_8i_s32f_convert_32f(float *out, const int8_t *in, const float scalar, unsigned int n)
{
    const float iScalar = 1.0 / scalar;
    for (unsigned int i = 0; i < n; i++) out[i] = ((float)in[i]) * iScalar;
}

/* volk_32f_s32f_convert_8i_generic: r = in * scalar; clamp to [INT8_MIN, INT8_MAX]; rintf */
void volk_32f_s32f_convert_8i(int8_t *out, const float *in, const float scalar, unsigned int n)
{
    const float min_val = -128.0f, max_val = 127.0f;
    for (unsigned int i = 0; i < n; i++) {
        float r = in[i] * scalar;
        if (r > max_val) r = max_val;
        else if (r < min_val) r = min_val;
        if (r != r) r = 0.0f;               /* NaN is UB in volk; pinned to 0 on both sides */
        out[i] = (int8_t)rintf(r);
    }
}

/* volk_32fc_s32fc_multiply_32fc, rounding of the x86 a_avx_fma / u_avx_fma kernels (what
 * volk's dispatcher picks on AVX2+FMA hosts): tmp = swap(a) * im(s); out = fmaddsub(a, re(s), tmp) */
void volk_32fc_s32fc_multiply_32fc(lv_32fc_t *out, const lv_32fc_t *in, const lv_32fc_t scalar, unsigned int n)
{
    const float sr = scalar.real(), si = scalar.imag();
    for (unsigned int i = 0; i < n; i++) {
        const float ar = in[i].real(), ai = in[i].imag();
        const float t0 = ai * si, t1 = ar * si;
        out[i] = lv_32fc_t(fmaf(ar, sr, -t0), fmaf(ai, sr, t1));
    }
}

/* volk_32fc_x2_conjugate_dot_prod_32fc_generic: two complex accumulators over point pairs */
void volk_32fc_x2_conjugate_dot_prod_32fc(lv_32fc_t *result, const lv_32fc_t *input, const lv_32fc_t *taps, unsigned int n)
{
    const float *in = (const float *)input, *tp = (const float *)taps;
    float sum0[2] = { 0, 0 }, sum1[2] = { 0, 0 };
    const unsigned int n_2 = n / 2;
    for (unsigned int i = 0; i < n_2; i++) {
        float a0 = in[0] * tp[0], a1 = in[1] * tp[1];
        float b0 = -in[0] * tp[1], b1 = in[1] * tp[0];
        float c0 = in[2] * tp[2], c1 = in[3] * tp[3];
        float d0 = -in[2] * tp[3], d1 = in[3] * tp[2];
        float e0 = a0 + a1, e1 = b0 + b1, e2 = c0 + c1, e3 = d0 + d1;
        sum0[0] += e0; sum0[1] += e1; sum1[0] += e2; sum1[1] += e3;
        in += 4; tp += 4;
    }
    float re = sum0[0] + sum1[0], im = sum0[1] + sum1[1];
    if (n & 1) {
        const lv_32fc_t a = input[n - 1], b = taps[n - 1];
        float t0 = a.real() * b.real(), t1 = a.imag() * b.imag();
        float t2 = a.imag() * b.real(), t3 = a.real() * b.imag();
        float u0 = t0 + t1, u1 = t2 - t3;
        re += u0; im += u1;
    }
    *result = lv_32fc_t(re, im);
}

/* volk_32f_x2_dot_prod_32f_generic */
void volk_32f_x2_dot_prod_32f(float *result, const float *in, const float *taps, unsigned int n)
{
    float acc = 0.0f;
    for (unsigned int i = 0; i < n; i++) { float p = in[i] * taps[i]; acc += p; }
    *result = acc;
}

/* volk_32fc_magnitude_squared_32f_generic: re*re + im*im */
void volk_32fc_magnitude_squared_32f(float *out, const lv_32fc_t *in, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++) {
        float a = in[i].real() * in[i].real(), b = in[i].imag() * in[i].imag();
        out[i] = a + b;
    }
}

/* volk_32fc_x2_multiply_conjugate_32fc_generic: a * conj(b) */
void volk_32fc_x2_multiply_conjugate_32fc(lv_32fc_t *out, const lv_32fc_t *a, const lv_32fc_t *b, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++) {
        const float ar = a[i].real(), ai = a[i].imag(), br = b[i].real(), bi = b[i].imag();
        float t0 = ar * br, t1 = ai * bi, t2 = ai * br, t3 = ar * bi;
        out[i] = lv_32fc_t(t0 + t1, t2 - t3);
    }
}

/* volk_32f_index_max_32u_generic: first index of the strict maximum */
void volk_32f_index_max_32u(uint32_t *target, const float *src, uint32_t n)
{
    if (n == 0) return;
    float mx = src[0]; uint32_t idx = 0;
    for (uint32_t i = 1; i < n; i++) if (src[i] > mx) { mx = src[i]; idx = i; }
    *target = idx;
}

/* ------------------------------------------------------------------------------- fftw ---- */
struct orc_shim_plan {
    int n, howmany, idist, odist, sign;
    fftwf_complex *in, *out;
};

extern "C" fftwf_complex *fftwf_alloc_complex(size_t n)
{
    return (fftwf_complex *)volk_malloc(sizeof(fftwf_complex) * n, 64);
}
extern "C" void fftwf_free(void *p) { free(p); }

extern "C" fftwf_plan fftwf_plan_many_dft(int rank, const int *n, int howmany, fftwf_complex *in, const int *,
                                          int istride, int idist, fftwf_complex *out, const int *,
                                          int ostride, int odist, int sign, unsigned)
{
    if (rank != 1 || istride != 1 || ostride != 1) return nullptr;      /* only what ccoherent.cc:87-93 asks */
    orc_shim_plan *p = new orc_shim_plan;
    p->n = n[0]; p->howmany = howmany; p->idist = idist; p->odist = odist; p->sign = sign;
    p->in = in; p->out = out;
    return p;
}
extern "C" void fftwf_execute_dft(const fftwf_plan p, fftwf_complex *in, fftwf_complex *out)
{
    for (int k = 0; k < p->howmany; k++)
        orc_fft((orc_cf32 *)(out + (size_t)k * p->odist), (const orc_cf32 *)(in + (size_t)k * p->idist), p->n, p->sign);
}
extern "C" void fftwf_execute(const fftwf_plan p) { fftwf_execute_dft(p, p->in, p->out); }
extern "C" void fftwf_destroy_plan(fftwf_plan p) { delete p; }

/* ---------------------------------------------------------------------------- librtlsdr ---- */
struct rtlsdr_dev {
    uint32_t index;
    std::atomic<bool> cancel{ false };
    std::mutex m;
    std::vector<float> corrections;          /* every rtlsdr_set_sample_freq_correction_f value */
};
static const uint32_t kFakeDevices = 256;

extern "C" uint32_t rtlsdr_get_device_count(void) { return kFakeDevices; }
extern "C" const char *rtlsdr_get_device_name(uint32_t) { return "oracle fake RTL-SDR"; }
extern "C" int rtlsdr_get_device_usb_strings(uint32_t index, char *manufact, char *product, char *serial)
{
    if (manufact) strcpy(manufact, "oracle");
    if (product) strcpy(product, "fake");
    if (serial) snprintf(serial, 256, "%u", index);
    return 0;
}
extern "C" int rtlsdr_get_index_by_serial(const char *serial) { return serial ? atoi(serial) : -1; }
extern "C" int rtlsdr_open(rtlsdr_dev_t **dev, uint32_t index)
{
    *dev = new rtlsdr_dev; (*dev)->index = index;
    return 0;
}
extern "C" int rtlsdr_close(rtlsdr_dev_t *dev) { delete dev; return 0; }
extern "C" int rtlsdr_set_center_freq(rtlsdr_dev_t *, uint32_t) { return 0; }
extern "C" int rtlsdr_set_sample_rate(rtlsdr_dev_t *, uint32_t) { return 0; }
extern "C" int rtlsdr_set_agc_mode(rtlsdr_dev_t *, int) { return 0; }
extern "C" int rtlsdr_set_tuner_gain(rtlsdr_dev_t *, int) { return 0; }
extern "C" int rtlsdr_set_tuner_gain_mode(rtlsdr_dev_t *, int) { return 0; }
extern "C" int rtlsdr_set_dithering(rtlsdr_dev_t *, int) { return 0; }
extern "C" int rtlsdr_set_sample_freq_correction_f(rtlsdr_dev_t *dev, float ppm)
{
    std::lock_guard<std::mutex> lock(dev->m);
    dev->corrections.push_back(ppm);
    return 0;
}
extern "C" int rtlsdr_reset_buffer(rtlsdr_dev_t *) { return 0; }
extern "C" int rtlsdr_read_async(rtlsdr_dev_t *dev, rtlsdr_read_async_cb_t, void *, uint32_t, uint32_t)
{
    /* no USB: the harness delivers blocks through csdrdevice::swapbuffer; park until cancelled */
    while (!dev->cancel.load()) std::this_thread::sleep_for(std::chrono::milliseconds(2));
    return 0;
}
extern "C" int rtlsdr_cancel_async(rtlsdr_dev_t *dev) { dev->cancel = true; return 0; }
