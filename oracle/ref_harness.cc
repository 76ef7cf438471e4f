/*
 * oracle/ref_harness.cc -- drives the reference's OWN code hardware-free.  TEST INFRASTRUCTURE
 * ONLY (checker and timed CPU baseline; never linked into or called by the product).
 *
 * oracle/Makefile compiles the reference's translation units (cdsp.cc, ccoherent.cc,
 * csdrdevice.cc, crtlsdr.cc, cpacketizer.cc, common.cc, ccontrol.cc) unmodified, from where they
 * lie under /root/reference, against the declaration shims in oracle/shim/, together with this
 * file, into oracle/_ref/libref_cdsp.so.  This file builds the same object graph as the
 * reference's main() (src/main.cc:226-268: crefsdr + N crtlsdr + cpacketize::init + crefnoise +
 * ccoherent) and feeds synthetic blocks through csdrdevice::swapbuffer -- the function
 * librtlsdr's callback calls (src/crtlsdr.cc:62-68,173-193) -- so that the reference's own
 * ccoherent::threadf / computelag / est_phasecorrect / phasecorrect / cpacketize::write run on
 * them, one block per push, and the bytes cpacketize::send() hands to ZMQ are captured.
 *
 * Determinism choices (documented deviations from a live run, none touching arithmetic):
 *  - the per-device ccontrol threads are never started (crtlsdr::start is not called): the
 *    harness sets each device's lag-request flag itself, so which blocks are correlated is an
 *    input, not a race against wall-clock;
 *  - every device is primed with `asyncbufn` buffers and the ring read pointers are aligned
 *    before the DSP thread starts, which is the state a live run reaches while FFTW plans
 *    (src/ccoherent.cc:87-93) -- all channels is_ready() and block-aligned;
 *  - cpacketize::send() is entered before the block is pushed, as in the live main loop
 *    (src/main.cc:277-279), so the header lands in the packet that is being filled.
 * The reference's private members are reached by compiling with g++ -fno-access-control
 * (oracle/Makefile); the reference sources themselves are compiled untouched.
 */
#include <atomic>
#include <chrono>
#include <complex>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <mutex>
#include <queue>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#include <unistd.h>

#include "ccoherent.h"
#include "csdrdevice.h"
#include "cpacketizer.h"
#include "cdsp.h"
#include "crefnoise.h"

extern zmq::socket_t *debugsocket;          /* src/cpacketizer.cc:44 */

namespace {
struct RefHarness {
    uint32_t C = 0, B = 0, nch = 0, asyncbufn = 8;
    int noheader = 0;
    crefsdr *ref = nullptr;
    lvector<csdrdevice *> devs;
    crefnoise *noise = nullptr;
    ccoherent *coh = nullptr;
    std::vector<std::vector<uint8_t>> ring;     /* [device][asyncbufn * B] stand-in for librtlsdr's buffers */
    uint64_t pushed = 0;
    csdrdevice *dev(uint32_t c) { return c == 0 ? (csdrdevice *)ref : devs[c - 1]; }
};
RefHarness *g_live = nullptr;                   /* cpacketize is all-static: one instance at a time */

void push_raw(RefHarness *h, const uint8_t *const *chan)
{
    const uint32_t slot = (uint32_t)(h->pushed % h->asyncbufn);
    for (uint32_t c = 0; c < h->nch; c++) {
        uint8_t *buf = h->ring[c].data() + (size_t)slot * h->B;
        if (chan) std::memcpy(buf, chan[c], h->B); else std::memset(buf, 0x80, h->B);
        h->dev(c)->swapbuffer(buf);             /* XORs in place (common.h:121), posts the buffer */
    }
    h->pushed++;
}
}  // namespace

extern "C" {

void *refh_create(uint32_t n_signal, uint32_t blocksize, int noheader, int nfft)
{
    if (g_live || n_signal == 0 || blocksize < 16 || (blocksize & (blocksize - 1))) return nullptr;
    RefHarness *h = new RefHarness;
    h->C = n_signal; h->B = blocksize; h->nch = n_signal + 1; h->noheader = noheader;
    const uint32_t fs = 2048000, fc = 1024000000u;                   /* src/main.cc:167 defaults */
    h->noise = new crefnoise("/dev/null");                           /* src/main.cc:185 opens /dev/ttyACM0 */
    h->ref = new crefsdr(h->asyncbufn, blocksize, fs, fc);
    h->ref->open(0u);
    for (uint32_t k = 0; k < n_signal; k++) {
        h->devs.push_back(new crtlsdr(h->asyncbufn, blocksize, fs, fc));
        h->devs.back()->open(k + 1);
    }
    cpacketize::globalseqn = 0;
    cpacketize::bufferfilled = false;
    cpacketize::do_exit = false;
    cpacketize::init("tcp://*:5555", noheader != 0, h->nch, blocksize); /* src/main.cc:258 */
    h->ring.resize(h->nch);
    for (auto &r : h->ring) r.assign((size_t)h->asyncbufn * blocksize, 0x80);
    /* prime: asyncbufn buffers per device, then align the rings (see header comment) */
    for (uint32_t c = 0; c < h->nch; c++) h->dev(c)->streaming = true;  /* crtlsdr::start / crefsdr::start */
    for (uint32_t i = 0; i < h->asyncbufn; i++) push_raw(h, nullptr);
    for (uint32_t c = 0; c < h->nch; c++) {
        crtlsdr *d = static_cast<crtlsdr *>(h->dev(c));
        d->newdata = 0;
        d->s8bit.rp = d->s8bit.wp;
        d->lagrequested = false;
    }
    h->coh = new ccoherent(h->ref, &h->devs, h->noise, nfft);
    h->coh->start();
    g_live = h;
    return h;
}

size_t refh_packet_len(void *hv)         /* bytes cpacketize::write fills: hdr0 + readcnt + columns */
{
    RefHarness *h = (RefHarness *)hv;
    return (h->noheader ? 0 : 16 + 4 * (size_t)h->nch) + (size_t)h->nch * h->B;
}
size_t refh_sent_len(void *) { return cpacketize::packetlen; }   /* what send() publishes (2x quirk) */

/* One block in lock-step through the reference's DSP thread.  chan: C+1 raw offset-binary
 * buffers (copied; the reference XORs its copy in place).  Outputs may be NULL. */
int refh_push_block(void *hv, const uint8_t *const *chan, const uint8_t *lag_request, int refnoise_enabled,
                    int8_t *packet, float *lag, float *mag, float *phase, uint8_t *measured, uint32_t *readcnt)
{
    RefHarness *h = (RefHarness *)hv;
    if (!h || h != g_live) return -1;
    h->noise->enabled = refnoise_enabled != 0;
    for (uint32_t c = 1; c < h->nch; c++) {
        csdrdevice *d = h->dev(c);
        d->lagready = false;
        if (lag_request && lag_request[c]) d->requestfft();          /* csdrdevice.h:128 */
        else d->lagrequested = false;
    }
    const uint32_t seq0 = cpacketize::globalseqn;
    std::thread sender([] { cpacketize::send(); });                  /* src/main.cc:277-279 */
    if (!h->noheader)
        while (*(volatile uint32_t *)&cpacketize::globalseqn == seq0) std::this_thread::yield();
    push_raw(h, chan);
    sender.join();
    const std::vector<uint8_t> &pk = cpacketize::socket->last;
    if (packet) std::memcpy(packet, pk.data(), refh_packet_len(h));
    for (uint32_t c = 0; c < h->nch; c++) {
        csdrdevice *d = h->dev(c);
        if (lag) lag[c] = c ? d->get_lagp()->lag : 0.0f;
        if (mag) mag[c] = c ? d->get_lagp()->mag : 0.0f;
        if (phase) {
            std::complex<float> p = c ? d->get_phasecorrect() : std::complex<float>(1.0f, 0.0f);
            phase[2 * c] = p.real(); phase[2 * c + 1] = p.imag();
        }
        if (measured) measured[c] = (c && d->lagready.load()) ? 1 : 0;
        if (readcnt) readcnt[c] = d->get_readcnt();
    }
    return 0;
}

/* contiguous batch, iq [n_blocks][C+1][B]; lag_request [n_blocks][C+1] or NULL (-> lag_mode for
 * every signal channel); refnoise [n_blocks] or NULL (enabled).  Outputs per block, may be NULL. */
int refh_run_batch(void *hv, const uint8_t *iq, uint32_t n_blocks, const uint8_t *lag_request, int lag_mode,
                   const uint8_t *refnoise, int8_t *packets, float *lag, float *mag, float *phase, uint8_t *measured)
{
    RefHarness *h = (RefHarness *)hv;
    if (!h || h != g_live) return -1;
    const size_t plen = refh_packet_len(h);
    std::vector<const uint8_t *> chan(h->nch);
    std::vector<uint8_t> flags(h->nch, (uint8_t)(lag_mode != 0));
    for (uint32_t b = 0; b < n_blocks; b++) {
        for (uint32_t c = 0; c < h->nch; c++) chan[c] = iq + ((size_t)b * h->nch + c) * h->B;
        const uint8_t *fl = lag_request ? lag_request + (size_t)b * h->nch : flags.data();
        int rc = refh_push_block(h, chan.data(), fl, refnoise ? refnoise[b] : 1,
                                 packets ? packets + (size_t)b * plen : nullptr,
                                 lag ? lag + (size_t)b * h->nch : nullptr, mag ? mag + (size_t)b * h->nch : nullptr,
                                 phase ? phase + (size_t)b * h->nch * 2 : nullptr,
                                 measured ? measured + (size_t)b * h->nch : nullptr, nullptr);
        if (rc) return rc;
    }
    return 0;
}

/* values recorded by the fake rtlsdr_set_sample_freq_correction_f are not needed here: the
 * controllers are not running.  Phase state for checkpoint tests: */
void refh_get_state(void *hv, float *phasecorr, float *phasecorrprev)
{
    RefHarness *h = (RefHarness *)hv;
    for (uint32_t c = 1; c < h->nch; c++) {
        phasecorr[2 * c] = h->dev(c)->phasecorr.real(); phasecorr[2 * c + 1] = h->dev(c)->phasecorr.imag();
        phasecorrprev[2 * c] = h->dev(c)->phasecorrprev.real(); phasecorrprev[2 * c + 1] = h->dev(c)->phasecorrprev.imag();
    }
}

void refh_destroy(void *hv)
{
    RefHarness *h = (RefHarness *)hv;
    if (!h || h != g_live) return;
    h->coh->request_exit();
    push_raw(h, nullptr);                       /* unblock refdev->read() so threadf sees do_exit */
    h->coh->join();
    delete h->coh;
    for (uint32_t c = 0; c < h->nch; c++) { h->dev(c)->streaming = false; h->dev(c)->close(); }
    for (uint32_t k = 0; k < h->C; k++) delete h->devs[k];
    delete h->ref;
    cpacketize::cleanup();
    delete h->noise;
    g_live = nullptr;
    delete h;
}

/* ---- live mode: the whole object graph free-running against the simulated dongles ---------
 * Everything main() starts (src/main.cc:226-279) is started for real: the per-device read threads
 * (crtlsdr::start -> rtlsdr_read_async -> asynch_callback -> swapbuffer), the per-device
 * controllers (ccontrol::threadf, the ppm law of src/ccontrol.cc:73-123), the DSP thread
 * (ccoherent::threadf) and a sender looping cpacketize::send().  The fake librtlsdr in
 * shim_impl.cc is a replay source whose delay moves with the correction the controller sets, so
 * the loop closes: lag measured by the path -> ccontrol -> resampler -> delay -> next lag.
 * Start order as in main(): devices, then ccoherent once the rings have filled.                */
extern "C" void orc_sim_configure(uint32_t n_dev, const double *delay0, const float *gain, const float *phase,
                                  double fs, uint64_t seed);
extern "C" void orc_sim_off(void);
extern "C" int orc_sim_get(uint32_t index, double *delay, float *corr, uint64_t *blocks, uint32_t *n_corrections);
extern "C" int orc_sim_kick(uint32_t index, double delta);

namespace {
struct LiveHarness {
    uint32_t C = 0, B = 0, nch = 0;
    crefsdr *ref = nullptr;
    lvector<csdrdevice *> devs;
    crefnoise *noise = nullptr;
    ccoherent *coh = nullptr;
    barrier *bar = nullptr;
    std::thread sender;
    std::atomic<bool> stop{ false };
    std::atomic<uint64_t> packets{ 0 };
    std::mutex pm;
    std::vector<uint8_t> last_packet;
    crtlsdr *dev(uint32_t c) { return c == 0 ? (crtlsdr *)ref : (crtlsdr *)devs[c - 1]; }
};
LiveHarness *g_livesim = nullptr;
}  // namespace

void *refh_live_create(uint32_t n_signal, uint32_t blocksize, uint32_t fs, const double *delay0, const float *gain,
                       const float *phase, uint64_t seed, int nfft)
{
    if (g_live || g_livesim || n_signal == 0 || blocksize < 16 || (blocksize & (blocksize - 1))) return nullptr;
    LiveHarness *h = new LiveHarness;
    h->C = n_signal; h->B = blocksize; h->nch = n_signal + 1;
    orc_sim_configure(h->nch, delay0, gain, phase, (double)fs, seed);
    const uint32_t fc = 1024000000u, asyncbufn = 8;
    h->noise = new crefnoise("/dev/null");
    h->noise->enabled = true;
    h->ref = new crefsdr(asyncbufn, blocksize, fs, fc);
    h->ref->open(0u);
    for (uint32_t k = 0; k < n_signal; k++) {
        h->devs.push_back(new crtlsdr(asyncbufn, blocksize, fs, fc));
        h->devs.back()->open(k + 1);
    }
    cpacketize::globalseqn = 0;
    cpacketize::bufferfilled = false;
    cpacketize::do_exit = false;
    cpacketize::init("tcp://*:5555", false, h->nch, blocksize);
    h->sender = std::thread([h] {
        while (!h->stop.load()) {
            cpacketize::send();                                      /* src/main.cc:277-279 */
            if (h->stop.load()) break;
            std::lock_guard<std::mutex> lock(h->pm);
            h->last_packet = cpacketize::socket->last;
            h->packets++;
        }
    });
    h->bar = new barrier(h->nch);
    h->ref->start(h->bar);
    for (uint32_t k = 0; k < n_signal; k++) h->devs[k]->start(h->bar);   /* read thread + controller */
    /* main() constructs ccoherent after the devices were started, and its FFTW_MEASURE planning
     * outlasts the first asyncbufn blocks: the DSP thread first runs with every ring full, which
     * is what keeps the signal rings block-aligned with the reference ring (threadf does not
     * consume() from a device that is not is_ready(), src/ccoherent.cc:262-281).  Made explicit: */
    for (uint32_t c = 0; c < h->nch; c++)
        while (h->dev(c)->get_readcnt() < asyncbufn) std::this_thread::sleep_for(std::chrono::milliseconds(1));
    h->coh = new ccoherent(h->ref, &h->devs, h->noise, nfft);
    h->coh->start();
    g_livesim = h;
    return h;
}

/* per channel (entry 0 = reference): last published lag, synced flag, the simulator's true delay,
 * the correction in force, how many corrections the controller has issued; returns packets sent */
uint64_t refh_live_poll(void *hv, float *lag, uint8_t *synced, double *delay, float *corr, uint32_t *n_corrections,
                        uint64_t *blocks)
{
    LiveHarness *h = (LiveHarness *)hv;
    if (!h || h != g_livesim) return 0;
    for (uint32_t c = 0; c < h->nch; c++) {
        if (lag) lag[c] = c ? h->dev(c)->get_lagp()->lag : 0.0f;
        if (synced) synced[c] = c ? (h->dev(c)->get_synchronized() ? 1 : 0) : 1;
        orc_sim_get(c, delay ? delay + c : nullptr, corr ? corr + c : nullptr, blocks ? blocks + c : nullptr,
                    n_corrections ? n_corrections + c : nullptr);
    }
    return h->packets.load();
}

/* an external delay step on one channel and, as the console's `sync` command does
 * (src/console.cc:285-289), the synced flag dropped so the controller re-acquires */
int refh_live_kick(void *hv, uint32_t chan, double delta)
{
    LiveHarness *h = (LiveHarness *)hv;
    if (!h || h != g_livesim || chan == 0 || chan >= h->nch) return -1;
    orc_sim_kick(chan, delta);
    /* the source works one block ahead: let the step reach the DSP thread before asking */
    const uint64_t seen = h->packets.load();
    while (h->packets.load() < seen + 3) std::this_thread::sleep_for(std::chrono::milliseconds(1));
    h->dev(chan)->set_synchronized(false);
    return 0;
}

size_t refh_live_last_packet(void *hv, uint8_t *out, size_t cap)
{
    LiveHarness *h = (LiveHarness *)hv;
    if (!h || h != g_livesim) return 0;
    std::lock_guard<std::mutex> lock(h->pm);
    const size_t n = h->last_packet.size() < cap ? h->last_packet.size() : cap;
    if (out && n) std::memcpy(out, h->last_packet.data(), n);
    return h->last_packet.size();
}

void refh_live_destroy(void *hv)
{
    LiveHarness *h = (LiveHarness *)hv;
    if (!h || h != g_livesim) return;
    /* controllers first, while blocks still flow: a controller parked in wait_synchronized() or
     * requestfftblocking() needs the DSP thread alive to come back and see do_exit */
    for (uint32_t k = 0; k < h->C; k++) {
        h->devs[k]->controller->request_exit();
        h->devs[k]->set_synchronized(false);
    }
    for (uint32_t k = 0; k < h->C; k++) h->devs[k]->controller->join();
    h->coh->request_exit();
    h->coh->join();                              /* sees do_exit at the next block */
    delete h->coh;
    h->stop = true;
    cpacketize::do_exit = true;
    cpacketize::cv.notify_all();
    h->sender.join();
    for (uint32_t c = 0; c < h->nch; c++) {      /* crtlsdr::stop() minus its raise(SIGUSR1) */
        rtlsdr_cancel_async(h->dev(c)->dev);
        h->dev(c)->streaming = false;
        h->dev(c)->cv.notify_all();
    }
    for (uint32_t c = 0; c < h->nch; c++) if (h->dev(c)->thread.joinable()) h->dev(c)->thread.join();
    for (uint32_t c = 0; c < h->nch; c++) h->dev(c)->close();
    for (uint32_t k = 0; k < h->C; k++) delete h->devs[k];
    delete h->ref;
    cpacketize::cleanup();
    delete h->noise;
    delete h->bar;
    orc_sim_off();
    g_livesim = nullptr;
    delete h;
}

/* ---- the reference's cdsp:: statics, called directly (src/cdsp.cc) ------------------------ */
void refh_cdsp_convtosigned(const uint8_t *in, uint8_t *out, int n) { cdsp::convtosigned(in, out, n); }
void refh_cdsp_convtofloat(float *out, const int8_t *s8, int n) { cdsp::convtofloat(out, s8, n); }
void refh_cdsp_convto8bit(int8_t *out, float *in, int n_complex)
{
    cdsp::convto8bit((std::complex<int8_t> *)out, (std::complex<float> *)in, n_complex);
}
void refh_cdsp_scalarmul(float *out, const float *in, float sr, float si, int n)
{
    cdsp::scalarmul((std::complex<float> *)out, (const std::complex<float> *)in, std::complex<float>(sr, si), n);
}
void refh_cdsp_conj_dotproduct(float *res2, const float *a, const float *b, int n)
{
    std::complex<float> r = cdsp::conj_dotproduct((const std::complex<float> *)a, (const std::complex<float> *)b, n);
    res2[0] = r.real(); res2[1] = r.imag();
}
float refh_cdsp_rms(const float *in, int n) { return cdsp::rms(in, n); }
float refh_cdsp_crestfactor(const float *in, int n) { return cdsp::crestfactor(in, n); }
void refh_cdsp_magsquared(float *out, const float *in, int n) { cdsp::magsquared(out, (const std::complex<float> *)in, n); }
void refh_cdsp_conjugatemul(float *out, float *a, float *b, int n)
{
    cdsp::conjugatemul((std::complex<float> *)out, (std::complex<float> *)a, (std::complex<float> *)b, n);
}
uint32_t refh_cdsp_indexofmax(float *in, int n) { return cdsp::indexofmax(in, n); }

}  // extern "C"
