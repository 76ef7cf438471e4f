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
#include <cmath>
#include <cstdio>
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
/*
 * Two modes.  Lock-step (default): rtlsdr_read_async parks and the harness pushes blocks through
 * csdrdevice::swapbuffer itself.  Replay/simulation (orc_sim_configure): every opened device is a
 * live source -- rtlsdr_read_async calls the reference's callback (crtlsdr::asynch_callback,
 * src/crtlsdr.cc:62-68) once per block period from a ring of buf_num buffers, the way librtlsdr
 * does -- with
 *     x_k[t] = g_k e^{j phi_k} r[t - round(d_k)] + w_k[t],   device 0: x_0 = r   (SURVEY 8d input)
 * r, w counter-hashed Gaussians (any t is addressable, so delays may be negative), and the
 * resampler model the reference's controller relies on (src/ccontrol.cc:96-116: after a
 * correction f held for a time t the delay has moved by f*fs*t samples):
 *     d_k  <-  d_k - f_k * L        once per block of L samples,
 * f_k the last value passed to rtlsdr_set_sample_freq_correction_f.  Pacing is wall-clock
 * (block period L/fs) because ccontrol sleeps on the wall clock.
 */
struct rtlsdr_dev {
    uint32_t index;
    std::atomic<bool> cancel{ false };
    std::mutex m;
    std::vector<float> corrections;          /* every rtlsdr_set_sample_freq_correction_f value */
    std::atomic<float> corr{ 0.0f };         /* the one in force */
    std::atomic<double> delay{ 0.0 };
    std::atomic<uint64_t> blocks{ 0 };
    float gain = 1.0f, phase = 0.0f;
};
static const uint32_t kFakeDevices = 256;

namespace {
struct OrcSim {
    std::mutex m;
    bool on = false;
    double fs = 0.0;
    uint64_t seed = 0;
    std::vector<double> delay0;
    std::vector<float> gain, phase;
    bool t0_set = false;
    std::chrono::steady_clock::time_point t0;
    rtlsdr_dev *open_dev[kFakeDevices] = {};
} g_sim;

inline uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
/* two unit-variance near-Gaussians (Irwin-Hall of four 16-bit uniforms each) per counter value */
inline void gauss_pair(uint64_t stream, int64_t t, float &a, float &b)
{
    const uint64_t h0 = splitmix64(stream ^ (uint64_t)t * 0xD1342543DE82EF95ull), h1 = splitmix64(h0);
    const float k = 1.0f / 37837.2f;         /* sqrt(4 * 65536^2 / 12) */
    a = ((float)((h0 & 0xffff) + ((h0 >> 16) & 0xffff) + ((h0 >> 32) & 0xffff) + (h0 >> 48)) - 131070.0f) * k;
    b = ((float)((h1 & 0xffff) + ((h1 >> 16) & 0xffff) + ((h1 >> 32) & 0xffff) + (h1 >> 48)) - 131070.0f) * k;
}
inline uint8_t quantise(float v)
{
    float q = std::nearbyint(v + 127.5f);
    return (uint8_t)(q < 0.0f ? 0.0f : (q > 255.0f ? 255.0f : q));
}
void sim_fill(rtlsdr_dev *dev, uint8_t *buf, uint32_t len, uint64_t n)
{
    const uint32_t L = len / 2;
    const int64_t d = (int64_t)std::llround(dev->delay.load());
    const uint64_t rs = g_sim.seed, ws = splitmix64(g_sim.seed + 1 + dev->index);
    const float cr = dev->gain * std::cos(dev->phase), ci = dev->gain * std::sin(dev->phase);
    for (uint32_t i = 0; i < L; i++) {
        const int64_t t = (int64_t)(n * L) + i;
        float rr, ri;
        gauss_pair(rs, t - d, rr, ri);
        rr *= 40.0f; ri *= 40.0f;
        float xr = rr, xi = ri;
        if (dev->index != 0) {
            float wr, wi;
            gauss_pair(ws, t, wr, wi);
            xr = cr * rr - ci * ri + 4.0f * wr;
            xi = cr * ri + ci * rr + 4.0f * wi;
        }
        buf[2 * i] = quantise(xr); buf[2 * i + 1] = quantise(xi);
    }
}
}  // namespace

/* n_dev devices (index 0 = reference); delay0/gain/phase per device (entry 0 ignored) */
extern "C" void orc_sim_configure(uint32_t n_dev, const double *delay0, const float *gain, const float *phase,
                                  double fs, uint64_t seed)
{
    std::lock_guard<std::mutex> lock(g_sim.m);
    g_sim.on = n_dev > 0; g_sim.fs = fs; g_sim.seed = seed; g_sim.t0_set = false;
    g_sim.delay0.assign(delay0, delay0 + n_dev);
    g_sim.gain.assign(gain, gain + n_dev);
    g_sim.phase.assign(phase, phase + n_dev);
}
extern "C" void orc_sim_off(void) { std::lock_guard<std::mutex> lock(g_sim.m); g_sim.on = false; }
extern "C" int orc_sim_get(uint32_t index, double *delay, float *corr, uint64_t *blocks, uint32_t *n_corrections)
{
    std::lock_guard<std::mutex> lock(g_sim.m);
    rtlsdr_dev *d = index < kFakeDevices ? g_sim.open_dev[index] : nullptr;
    if (!d) return -1;
    if (delay) *delay = d->delay.load();
    if (corr) *corr = d->corr.load();
    if (blocks) *blocks = d->blocks.load();
    if (n_corrections) { std::lock_guard<std::mutex> l2(d->m); *n_corrections = (uint32_t)d->corrections.size(); }
    return 0;
}
/* a delay step injected from outside (cable swap / dropped USB transfer), for re-sync tests */
extern "C" int orc_sim_kick(uint32_t index, double delta)
{
    std::lock_guard<std::mutex> lock(g_sim.m);
    rtlsdr_dev *d = index < kFakeDevices ? g_sim.open_dev[index] : nullptr;
    if (!d) return -1;
    d->delay.store(d->delay.load() + delta);
    return 0;
}

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
    rtlsdr_dev *d = new rtlsdr_dev; d->index = index;
    std::lock_guard<std::mutex> lock(g_sim.m);
    if (g_sim.on && index < g_sim.delay0.size()) {
        d->delay = index ? g_sim.delay0[index] : 0.0;
        d->gain = g_sim.gain[index]; d->phase = g_sim.phase[index];
    }
    if (index < kFakeDevices) g_sim.open_dev[index] = d;
    *dev = d;
    return 0;
}
extern "C" int rtlsdr_close(rtlsdr_dev_t *dev)
{
    {
        std::lock_guard<std::mutex> lock(g_sim.m);
        if (dev->index < kFakeDevices && g_sim.open_dev[dev->index] == dev) g_sim.open_dev[dev->index] = nullptr;
    }
    delete dev;
    return 0;
}
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
    dev->corr = ppm;
    return 0;
}
extern "C" int rtlsdr_reset_buffer(rtlsdr_dev_t *) { return 0; }
extern "C" int rtlsdr_read_async(rtlsdr_dev_t *dev, rtlsdr_read_async_cb_t cb, void *ctx, uint32_t buf_num, uint32_t buf_len)
{
    bool live; double fs; std::chrono::steady_clock::time_point t0;
    {
        std::lock_guard<std::mutex> lock(g_sim.m);
        live = g_sim.on && dev->index < g_sim.delay0.size() && cb && buf_num && buf_len;
        if (live && !g_sim.t0_set) { g_sim.t0 = std::chrono::steady_clock::now(); g_sim.t0_set = true; }
        fs = g_sim.fs; t0 = g_sim.t0;
    }
    if (!live) {
        /* no USB: the harness delivers blocks through csdrdevice::swapbuffer; park until cancelled */
        while (!dev->cancel.load()) std::this_thread::sleep_for(std::chrono::milliseconds(2));
        return 0;
    }
    std::vector<uint8_t> ring((size_t)buf_num * buf_len);
    const uint32_t L = buf_len / 2;
    const double period = (double)L / fs;
    for (uint64_t n = 0; !dev->cancel.load(); n++) {
        if (dev->index) dev->delay.store(dev->delay.load() - (double)dev->corr.load() * (double)L);
        uint8_t *buf = ring.data() + (size_t)(n % buf_num) * buf_len;
        sim_fill(dev, buf, buf_len, n);
        std::this_thread::sleep_until(t0 + std::chrono::duration_cast<std::chrono::steady_clock::duration>(
                                               std::chrono::duration<double>((double)(n + 1) * period)));
        if (dev->cancel.load()) break;
        cb(buf, buf_len, ctx);
        dev->blocks = n + 1;
    }
    return 0;
}
extern "C" int rtlsdr_cancel_async(rtlsdr_dev_t *dev) { dev->cancel = true; return 0; }
