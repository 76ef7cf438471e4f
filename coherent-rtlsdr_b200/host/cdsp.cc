// host/cdsp.cc -- drop-in replacement for the reference's src/cdsp.cc.
//
// Compiled against the reference's OWN, unchanged include/cdsp.h (class cdsp, 17 statics,
// include/cdsp.h:36-71): same names, same argument meaning, same return values.  Every body
// forwards to the matching cohcuda_cdsp_* entry point of libcoherent_cuda.so
// (include/coherent_cuda.h), i.e. to one sm_100a kernel; nothing is computed on the host and
// there is no volk/fftw fallback -- if the library or the GPU is missing the first call aborts
// with a message (the reference's DSP path has no error returns on these functions either,
// src/cdsp.cc:21-146).
//
// This is the slow, call-compatible path for code that still calls cdsp:: per buffer
// (csdrdevice::est_phasecorrect, cpacketize::write(..complex<float>*)); the per-block fast path
// is host/ccoherent.cc, which hands whole blocks to cohcuda_submit_block.
//
// fft_scheme: the reference's typedef stays (fftwf_plan, an opaque pointer, include/cdsp.h:31).
// Plans handed to cdsp::fft must come from cohcuda_host_make_fft_scheme() below -- the
// replacement for fftwf_plan_many_dft(rank 1, n, howmany, stride 1, dist n, sign)
// (src/ccoherent.cc:87-93); a plan made by the real fftw cannot be executed here.
#include "cdsp.h"

#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <utility>
#include <vector>

#include "coherent_cuda.h"
#include "cohcuda_host.h"

// what an fft_scheme points to on this path (fftw3.h keeps fftwf_plan an opaque pointer)
struct cohcuda_fft_scheme {
    int n, howmany, sign;
    std::complex<float> *in, *out;
};

namespace {
cohcuda_t *g_prim = nullptr;
std::mutex g_prim_mtx;
// a libcoherent_cuda handle takes one caller at a time; cdsp:: statics are reached from every
// librtlsdr callback thread (cbuffer::setbufferptr -> cdsp::convtosigned, include/common.h:121)
std::mutex g_call_mtx;

[[noreturn]] void die(const char *what, cohcuda_t *h)
{
    std::fprintf(stderr, "cdsp (libcoherent_cuda): %s: %s\n", what, cohcuda_last_error(h));
    std::abort();
}
#define PRIM(call) do { cohcuda_t *h__ = cohcuda_host_prim_handle(); std::lock_guard<std::mutex> lk__(g_call_mtx); \
                        if ((call) != COHCUDA_OK) die(#call, h__); } while (0)
}  // namespace

extern "C" cohcuda_t *cohcuda_host_prim_handle(void)
{
    std::lock_guard<std::mutex> lock(g_prim_mtx);
    if (!g_prim) {
        cohcuda_config cfg{};
        cfg.n_signal = 1; cfg.blocksize = 64; cfg.alpha = 0.5f;
        const char *dev = std::getenv("COHCUDA_DEVICE");
        cfg.device = dev ? std::atoi(dev) : 0;
        if (cohcuda_create(&cfg, &g_prim) != COHCUDA_OK) die("cohcuda_create", nullptr);
    }
    return g_prim;
}

extern "C" void *cohcuda_host_make_fft_scheme(int n, int howmany, int sign, void *in, void *out)
{
    cohcuda_fft_scheme *p = new cohcuda_fft_scheme;
    p->n = n; p->howmany = howmany; p->sign = sign;
    p->in = (std::complex<float> *)in; p->out = (std::complex<float> *)out;
    return p;
}
extern "C" void cohcuda_host_destroy_fft_scheme(void *scheme) { delete (cohcuda_fft_scheme *)scheme; }

#define H cohcuda_host_prim_handle()

// src/cdsp.cc:21-34 (n bytes, multiple of 8; `out` is written despite its const qualifier)
//
// The one hot caller is cbuffer::setbufferptr (include/common.h:121): every librtlsdr callback thread
// converts its 16 KiB USB buffer IN PLACE, 250 times a second per dongle.  Routing that through the GPU
// would be a synchronised H2D + kernel + D2H round trip under one mutex per USB buffer (5500 per second
// for URA21).  The drop-in therefore DEFERS the in-place form: the bytes stay offset-binary, the DSP
// thread hands them to cohcuda_submit_block with input_is_signed = 0 and every kernel applies ^0x80 in
// its load; column 0 of the packet is then taken from the library's packet, not from the
// raw buffer.  Out-of-place calls (a caller that wants the converted bytes) still run the kernel.
// COHCUDA_HOST_XOR_ON_CALL=1 restores the convert-at-call behaviour for in-place calls too.
//
// The primitive's contract is kept for everyone who reads such a buffer THROUGH cdsp:: -- a deferred
// range is remembered, and cdsp::convtofloat (the only cdsp:: static that reads int8) applies the
// pending ^0x80 on the device when its source lies inside one.  That is the path of the reference's
// console `phase` command (src/console.cc:295-299: d->convtofloat() then est_phasecorrect() on the ring
// buffers).  Code that reads the raw bytes of a deferred buffer directly still sees offset-binary
// (cohcuda_host_xor_pending() tells); the replaced ccoherent thread is the one such reader.
extern "C" int cohcuda_host_xor_deferred(void)
{
    static const int deferred = [] { const char *e = std::getenv("COHCUDA_HOST_XOR_ON_CALL"); return (e && std::atoi(e)) ? 0 : 1; }();
    return deferred;
}
namespace {
// byte ranges whose in-place XOR was deferred (the librtlsdr ring: asyncbufn = 8 buffers per dongle, fixed
// addresses for the life of the device, include/common.h:114-122) -- a few hundred entries at most
std::mutex g_pending_mtx;
std::vector<std::pair<const uint8_t *, size_t>> g_pending;
bool pending_covers(const void *p, size_t n)
{
    const uint8_t *b = static_cast<const uint8_t *>(p);
    std::lock_guard<std::mutex> lk(g_pending_mtx);
    for (const auto &r : g_pending)
        if (b >= r.first && b + n <= r.first + r.second) return true;
    return false;
}
void pending_add(const uint8_t *p, size_t n)
{
    std::lock_guard<std::mutex> lk(g_pending_mtx);
    for (auto &r : g_pending)
        if (r.first == p) { if (n > r.second) r.second = n; return; }
    g_pending.emplace_back(p, n);
}
// int8 -> float of a source that may still be offset-binary: the pending XOR runs as its own kernel first
void convtofloat_any(float *out, const int8_t *s8bit, int n)
{
    if (n > 0 && cohcuda_host_xor_deferred() && pending_covers(s8bit, (size_t)n)) {
        const int n8 = (n + 7) & ~7;                                  // the XOR kernel works on 8-byte words
        std::vector<uint8_t> raw((size_t)n8, 0x80), sgn((size_t)n8);
        std::copy(reinterpret_cast<const uint8_t *>(s8bit), reinterpret_cast<const uint8_t *>(s8bit) + n, raw.begin());
        PRIM(cohcuda_cdsp_convtosigned(H, raw.data(), sgn.data(), n8));
        PRIM(cohcuda_cdsp_convtofloat(H, out, reinterpret_cast<const int8_t *>(sgn.data()), n));
        return;
    }
    PRIM(cohcuda_cdsp_convtofloat(H, out, s8bit, n));
}
}  // namespace

extern "C" int cohcuda_host_xor_pending(const void *p, size_t n) { return pending_covers(p, n) ? 1 : 0; }
extern "C" void cohcuda_host_xor_forget(const void *p)
{
    std::lock_guard<std::mutex> lk(g_pending_mtx);
    for (size_t i = 0; i < g_pending.size(); i++)
        if (g_pending[i].first == p) { g_pending.erase(g_pending.begin() + i); return; }
}

void cdsp::convtosigned(const uint8_t *in, const uint8_t *out, int n)
{
    if (in == out && cohcuda_host_xor_deferred()) { if (n > 0) pending_add(in, (size_t)n); return; }
    PRIM(cohcuda_cdsp_convtosigned(H, in, const_cast<uint8_t *>(out), n));
}

// src/cdsp.cc:36-39
const float *cdsp::convtofloat(const float *out, const int8_t *s8bit, int n)
{
    convtofloat_any(const_cast<float *>(out), s8bit, n);
    return out;
}

// src/cdsp.cc:41-44 (n = number of int8 values = 2 x complex samples)
const std::complex<float> *cdsp::convtofloat(const std::complex<float> *out, const int8_t *s8bit, int n)
{
    convtofloat_any(reinterpret_cast<float *>(const_cast<std::complex<float> *>(out)), s8bit, n);
    return out;
}

// src/cdsp.cc:46-49
const std::complex<float> *cdsp::scalarmul(const std::complex<float> *out, const std::complex<float> *in,
                                           const std::complex<float> scalar_in, int n)
{
    PRIM(cohcuda_cdsp_scalarmul(H, reinterpret_cast<float *>(const_cast<std::complex<float> *>(out)),
                                reinterpret_cast<const float *>(in), scalar_in.real(), scalar_in.imag(), n));
    return out;
}

// src/cdsp.cc:51-54 (n complex samples -> 2n int8)
const std::complex<int8_t> *cdsp::convto8bit(std::complex<int8_t> *out, std::complex<float> *in, int n)
{
    PRIM(cohcuda_cdsp_convto8bit(H, reinterpret_cast<int8_t *>(out), reinterpret_cast<const float *>(in), n));
    return out;
}

// src/cdsp.cc:61-66: sum a[i] * conj(b[i])
const std::complex<float> cdsp::conj_dotproduct(const std::complex<float> *a, const std::complex<float> *b, int n)
{
    float r[2] = { 0.0f, 0.0f };
    PRIM(cohcuda_cdsp_conj_dotproduct(H, r, reinterpret_cast<const float *>(a), reinterpret_cast<const float *>(b), n));
    return std::complex<float>(r[0], r[1]);
}

// src/cdsp.cc:68-73
const float cdsp::rms(const float *in, int n)
{
    float r = 0.0f;
    PRIM(cohcuda_cdsp_rms(H, &r, in, n));
    return r;
}

// src/cdsp.cc:75-78: sqrt(re(sum s conj s) / n)
const float cdsp::rms(const std::complex<float> *s, int n)
{
    std::complex<float> res = conj_dotproduct(s, s, n);
    return std::sqrt(res.real() / n);
}

// src/cdsp.cc:80-83
const float cdsp::crestfactor(const float *in, float peak, int n)
{
    float r = rms(in, n);
    return (peak / r);
}

// src/cdsp.cc:85-88: a stub in the reference ("FIXXXX"), returns 0
const float cdsp::PAPR(const std::complex<float> *, const std::complex<float> *, int) { return 0; }

// src/cdsp.cc:90-98
const float cdsp::crestfactor(const float *in, int n)
{
    uint32_t idx = 0;
    PRIM(cohcuda_cdsp_indexofmax(H, &idx, in, n));
    float m = in[idx];
    float r = rms(in, n);
    return (m / r);
}

// src/cdsp.cc:100-103
const float *cdsp::magsquared(float *out, const std::complex<float> *in, int n)
{
    PRIM(cohcuda_cdsp_magsquared(H, out, reinterpret_cast<const float *>(in), n));
    return out;
}

// src/cdsp.cc:105-108: in1 .* conj(in2)
const std::complex<float> *cdsp::conjugatemul(std::complex<float> *out, std::complex<float> *in1,
                                              std::complex<float> *in2, int n)
{
    PRIM(cohcuda_cdsp_conjugatemul(H, reinterpret_cast<float *>(out), reinterpret_cast<const float *>(in1),
                                   reinterpret_cast<const float *>(in2), n));
    return out;
}

// src/cdsp.cc:110-120: howmany transforms of length n, unnormalised, new-array execute
const std::complex<float> *cdsp::fft(std::complex<float> *out, std::complex<float> *in, fft_scheme *scheme)
{
    const cohcuda_fft_scheme *p = reinterpret_cast<const cohcuda_fft_scheme *>(*scheme);
    PRIM(cohcuda_cdsp_fft(H, reinterpret_cast<float *>(out), reinterpret_cast<const float *>(in), p->n, p->howmany, p->sign));
    return out;
}

// src/cdsp.cc:122-133: execute on the arrays the plan was made with; returns NULL like the reference
const std::complex<float> *cdsp::fft(fft_scheme *scheme)
{
    const cohcuda_fft_scheme *p = reinterpret_cast<const cohcuda_fft_scheme *>(*scheme);
    PRIM(cohcuda_cdsp_fft(H, reinterpret_cast<float *>(p->out), reinterpret_cast<const float *>(p->in), p->n, p->howmany, p->sign));
    return NULL;
}

// src/cdsp.cc:135-139: first index of the maximum
const uint32_t cdsp::indexofmax(float *in, int n)
{
    uint32_t idx = 0;
    PRIM(cohcuda_cdsp_indexofmax(H, &idx, in, n));
    return idx;
}

// src/cdsp.cc:141-146: |in|^2 into out, then its arg-max
const uint32_t cdsp::indexofmax(float *out, std::complex<float> *in, int n)
{
    uint32_t idx = 0;
    PRIM(cohcuda_cdsp_magsquared(H, out, reinterpret_cast<const float *>(in), n));
    PRIM(cohcuda_cdsp_indexofmax(H, &idx, out, n));
    return idx;
}
