// host/ccoherent.cc -- drop-in replacement for the reference's src/ccoherent.cc.
//
// Compiled against the reference's OWN, unchanged include/ccoherent.h: same class, same public
// surface (ctor(crefsdr*, lvector<csdrdevice*>*, crefnoise*, nfft), start / request_exit / join,
// lagqueuesize / clearlagqueue / queuelag / computelag, include/ccoherent.h:61-72), same single
// DSP thread (src/ccoherent.cc:245-298).  The difference is where the arithmetic runs: instead
// of convtofloat -> queuelag memcpy -> est_phasecorrect -> phasecorrect -> convto8bit per channel
// and an fftw batch per block on this thread, each block's raw buffers are handed to
// libcoherent_cuda (cohcuda_submit_block / cohcuda_wait_block, include/coherent_cuda.h) and the
// results are published through the reference's own, untouched interfaces:
//   packet columns   -> cpacketize::write(ch, readcnt, int8_t*)      (src/cpacketizer.cc:137-156)
//   phase factors    -> cpacketize::writedebug                        (src/cpacketizer.cc:131-134)
//   lag / magnitude  -> csdrdevice::set_lag (wakes the ccontrol thread, include/csdrdevice.h:138-151)
//   hand-over        -> cpacketize::notifysend                        (src/cpacketizer.cc:174-185)
// cpacketizer.cc, csdrdevice.cc, crtlsdr.cc, ccontrol.cc, common.cc and main.cc stay as they are.
//
// The FFT members of the class (fftscheme, sfft, sifft, sfloat, sconv, smagsqr) are unused on
// this path and left null; the engine handle lives in a side table keyed by `this` because the
// header is not modified.
//
// One private-access need: after a block the device objects' phasecorr / phasecorrprev must hold
// what the GPU computed (console `phase` and get_phasecorrect() read them,
// src/console.cc:294-332).  They are protected members of csdrdevice, so the maintainer adds
// `friend class ccoherent;` to class csdrdevice (one line) -- or builds this file
// with -fno-access-control, which is what the hardware-free harness in oracle/ does.
#include "ccoherent.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <map>
#include <mutex>
#include <vector>

#include "coherent_cuda.h"
#include "cohcuda_host.h"

namespace {

struct Engine {
    cohcuda_t *h = nullptr;
    uint32_t n_signal = 0, B = 0;
    std::vector<const uint8_t *> chan;          // [0] = reference, 1.. = signal
    std::vector<uint32_t> readcnt;
    std::vector<uint8_t> request, ready;
    std::vector<const void *> pinned;           // librtlsdr buffers already registered
    cohcuda_result res{};                       // result of the block in flight
    bool have_result = false;
    // which device sat in which engine channel, and whether it streamed, in the previous block: when that
    // changes (a dongle starts or stops, console add / del) the engine's per-channel state no longer
    // belongs to the device in that slot and is re-seeded from the devices' own phasecorr / phasecorrprev
    std::vector<std::pair<const void *, uint8_t>> slots;
    // error bookkeeping of the per-block path (see drop_block below)
    uint64_t dropped = 0;                       // blocks consumed without a packet because the library call failed
    uint32_t consecutive = 0;                   // ... in a row; kRebuildAfter of them rebuild the engine once
    bool rebuilt = false, pin_warned = false;
};
constexpr uint32_t kRebuildAfter = 50;
std::atomic<int> g_inject_failures{0};          // test hook: the next n submits are treated as failed

std::map<const ccoherent *, Engine *> g_engines;
std::mutex g_engines_mtx;

Engine *engine_of(const ccoherent *c)
{
    std::lock_guard<std::mutex> lock(g_engines_mtx);
    auto it = g_engines.find(c);
    return it == g_engines.end() ? nullptr : it->second;
}

[[noreturn]] void die(const char *what, cohcuda_t *h)
{
    // the reference's DSP thread has no error channel either; without the GPU there is no path
    std::fprintf(stderr, "ccoherent (libcoherent_cuda): %s: %s\n", what, cohcuda_last_error(h));
    std::abort();
}

// (re)build the engine for the current number of signal devices; per-channel state of the
// channels that already existed is carried over (console add/del, src/console.cc)
void size_engine(Engine *e, uint32_t n_signal, uint32_t blocksize)
{
    if (e->h && e->n_signal == n_signal) return;
    std::vector<cohcuda_chan_state> keep;
    if (e->h) {
        keep.resize(e->n_signal + 1);
        if (cohcuda_get_state(e->h, keep.data(), e->n_signal + 1) != COHCUDA_OK) die("cohcuda_get_state", e->h);
        cohcuda_destroy(e->h);
        e->h = nullptr; e->pinned.clear();
    }
    cohcuda_config cfg{};
    cfg.n_signal = n_signal;
    cfg.blocksize = blocksize;
    cfg.max_lag_slots = 0;              // the nfft-1 cap is applied by queuelag() below, as in the reference
    cfg.alpha = 0.5f;                   // src/csdrdevice.cc:60
    cfg.noheader = 1;                   // columns only: header + readcnt are cpacketize's job
    cfg.apply_shift = 0;                // the reference removes delay with the dongles' resamplers (src/ccontrol.cc)
    const char *dev = std::getenv("COHCUDA_DEVICE");
    cfg.device = dev ? std::atoi(dev) : 0;
    cfg.max_batch_blocks = 1;
    if (cohcuda_create(&cfg, &e->h) != COHCUDA_OK) die("cohcuda_create", nullptr);
    if (!keep.empty()) {
        uint32_t n = (uint32_t)keep.size() < n_signal + 1 ? (uint32_t)keep.size() : n_signal + 1;
        if (cohcuda_set_state(e->h, keep.data(), n) != COHCUDA_OK) die("cohcuda_set_state", e->h);
    }
    e->n_signal = n_signal; e->B = blocksize;
    e->chan.assign(n_signal + 1, nullptr);
    e->readcnt.assign(n_signal + 1, 0);
    e->request.assign(n_signal + 1, 0);
    e->ready.assign(n_signal + 1, 0);
    e->have_result = false;
}

// librtlsdr owns its USB buffers and recycles the same asyncbufn pointers per device
// (include/common.h:118-120): pin each one the first time it is seen.  Best effort -- an
// unpinned buffer still works, its copy is just staged by the driver.
void pin_once(Engine *e, const void *p, size_t bytes)
{
    for (const void *q : e->pinned) if (q == p) return;
    e->pinned.push_back(p);
    if (cohcuda_register_host(e->h, const_cast<void *>(p), bytes) != COHCUDA_OK && !e->pin_warned) {
        e->pin_warned = true;                   // said once: the path still works, only slower
        std::fprintf(stderr, "ccoherent (libcoherent_cuda): a USB buffer could not be pinned (%s); its copies are staged by the driver\n",
                     cohcuda_last_error(e->h));
    }
}

// A failed library call on the per-block path must not take a receiver down that has been running for days
// next to hot-plugged USB hardware.  The block is dropped: every device's buffer is consumed, nothing is
// written or sent, and the clients see the gap in readcnt / globalseqn exactly as they do when the reference
// loses USB transfers (src/cpacketizer.cc:137-156, 113).  kRebuildAfter failures in a row rebuild the engine
// once (a sticky CUDA error does not heal by itself); if that does not help either there is no path left --
// no CPU fallback exists by design -- and the process stops with the library's message.
void drop_block(ccoherent *ctx, Engine *e, const char *what)
{
    e->dropped++; e->consecutive++;
    if (e->consecutive == 1 || e->consecutive % 1000 == 0)
        std::fprintf(stderr, "ccoherent (libcoherent_cuda): %s: %s -- block dropped (%llu so far)\n", what,
                     e->h ? cohcuda_last_error(e->h) : "no engine", (unsigned long long)e->dropped);
    ctx->refdev->consume();
    uint32_t k = 1;
    for (auto *d : *ctx->devices) { if (k < e->ready.size() && e->ready[k]) d->consume(); k++; }
    if (e->consecutive >= kRebuildAfter) {
        if (e->rebuilt) die(what, e->h);
        std::fprintf(stderr, "ccoherent (libcoherent_cuda): %u failures in a row, rebuilding the engine\n", e->consecutive);
        std::vector<cohcuda_chan_state> keep(e->n_signal + 1);
        const bool have = e->h && cohcuda_get_state(e->h, keep.data(), e->n_signal + 1) == COHCUDA_OK;
        if (e->h) cohcuda_destroy(e->h);
        e->h = nullptr; e->pinned.clear(); e->slots.clear();
        const uint32_t n = e->n_signal, B = e->B;
        e->n_signal = 0;
        size_engine(e, n, B);                   // dies if the device is gone for good
        if (have) (void)cohcuda_set_state(e->h, keep.data(), n + 1);
        e->rebuilt = true; e->consecutive = 0;
    }
}

}  // namespace

extern "C" cohcuda_t *cohcuda_host_engine(const void *inst)
{
    Engine *e = engine_of((const ccoherent *)inst);
    return e ? e->h : nullptr;
}
extern "C" uint64_t cohcuda_host_dropped_blocks(const void *inst)
{
    std::lock_guard<std::mutex> lock(g_engines_mtx);
    if (inst) { auto it = g_engines.find((const ccoherent *)inst); return it == g_engines.end() ? 0 : it->second->dropped; }
    uint64_t n = 0;
    for (auto &kv : g_engines) n += kv.second->dropped;
    return n;
}
extern "C" void cohcuda_host_inject_failures(int n) { g_inject_failures.store(n > 0 ? n : 0); }

// src/ccoherent.cc:32-95 -- buffers and FFT plans become one library handle
ccoherent::ccoherent(crefsdr *refdev_, lvector<csdrdevice *> *devvec_, crefnoise *refnoise_, int nfft_)
{
    refdev = refdev_;
    devices = devvec_;
    refnoise = refnoise_;
    do_exit = false;
    nfft = nfft_;
    blocksize = refdev->get_blocksize();
    packetize = nullptr;
    fftscheme = ifftscheme = nullptr;
    sifft = sfft = sfloat = sconv = nullptr;
    smagsqr = nullptr;
    mb = 0;
    Engine *e = new Engine;
    size_engine(e, (uint32_t)devices->size() ? (uint32_t)devices->size() : 1u, (uint32_t)blocksize);
    std::lock_guard<std::mutex> lock(g_engines_mtx);
    g_engines[this] = e;
}

// src/ccoherent.cc:97-113
ccoherent::~ccoherent()
{
    if (thread.joinable()) thread.join();
    Engine *e = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_engines_mtx);
        auto it = g_engines.find(this);
        if (it != g_engines.end()) { e = it->second; g_engines.erase(it); }
    }
    if (e) { if (e->h) cohcuda_destroy(e->h); delete e; }
}

// src/ccoherent.cc:115-121
void ccoherent::clearlagqueue() { lagqueue.clear(); }
size_t ccoherent::lagqueuesize() { return lagqueue.size(); }

// src/ccoherent.cc:123-142: the 128 KiB memcpy into the FFT batch is gone; what remains is the
// bookkeeping -- slot 0 is the reference, at most nfft-1 signal channels per block (:124)
void ccoherent::queuelag(csdrdevice *d)
{
    if (lagqueue.size() < (size_t)nfft) lagqueue.push_back(d);
}
void ccoherent::queuelag(csdrdevice *d, bool) { queuelag(d); }   // declared in the header, unused by the reference

// src/ccoherent.cc:154-239: the correlation already ran on the GPU as part of the block; this
// publishes lag = idx - B/2 and mag = sqrt(peak/(B/2)) (:203-233) to the queued devices
void ccoherent::computelag()
{
    Engine *e = engine_of(this);
    if (e && e->have_result) {
        for (size_t q = 1; q < lagqueue.size(); q++) {
            csdrdevice *d = lagqueue[q];
            uint32_t k = 1;
            for (auto *dev : *devices) { if (dev == d) break; k++; }
            if (k <= e->n_signal && e->res.measured[k]) d->set_lag(e->res.lag[k], e->res.mag[k]);
        }
    }
    lagqueue.clear();                                                      // :235
}

// src/ccoherent.cc:245-294, same schedule, one library call per block
void ccoherent::threadf(ccoherent *ctx)
{
    Engine *e = engine_of(ctx);
    while (!ctx->do_exit) {
        ctx->clearlagqueue();
        int8_t *refsptr = ctx->refdev->read();
        if (ctx->do_exit) break;
        size_engine(e, (uint32_t)ctx->devices->size() ? (uint32_t)ctx->devices->size() : 1u, (uint32_t)ctx->blocksize);
        const uint32_t B = e->B;

        e->chan[0] = (const uint8_t *)refsptr;
        e->readcnt[0] = ctx->refdev->get_readcntbuf();
        e->request[0] = 0; e->ready[0] = 1;
        // A lag is only measured on blocks taken from the head of the rings.  While the thread works
        // off a backlog (start-up: the rings fill while the engine is created, as they do in the
        // reference while FFTW plans) a slot can be overwritten between the reads of two devices,
        // and a correlation across ring generations returns a meaningless peak that sends the
        // controller slewing the wrong way for seconds.  The request stays pending until then.
        const bool ref_at_head = ctx->refdev->newdata.load() == 0;
        pin_once(e, refsptr, B);
        ctx->queuelag(ctx->refdev);                                        // slot 0 (:253)

        uint32_t k = 1;
        for (auto *d : *ctx->devices) {
            int8_t *ptr = d->read();
            const bool ready = d->is_ready();
            e->ready[k] = ready;
            // a channel that is not streaming yet is skipped by the reference (:264); the engine
            // has a fixed channel count, so it is fed the reference buffer and its column dropped
            e->chan[k] = ready ? (const uint8_t *)ptr : (const uint8_t *)refsptr;
            e->readcnt[k] = d->get_readcntbuf();
            e->request[k] = 0;
            if (ready) {
                pin_once(e, ptr, B);
                if (d->is_lagrequested() && ref_at_head && d->newdata.load() == 0) {   // :266-267
                    const size_t before = ctx->lagqueuesize();
                    ctx->queuelag(d);
                    e->request[k] = ctx->lagqueuesize() > before;
                }
            }
            k++;
        }
        if (ctx->devices->size() == 0) { e->chan[1] = e->chan[0]; e->request[1] = 0; e->ready[1] = 0; }
        {
            // csdrdevice::phasecorr is the truth (include/csdrdevice.h:74-76; a skipped device keeps it untouched,
            // src/ccoherent.cc:264): the engine's copy follows it whenever the slot assignment changed
            std::vector<std::pair<const void *, uint8_t>> now;
            uint32_t j = 1;
            for (auto *d : *ctx->devices) { now.emplace_back((const void *)d, e->ready[j]); j++; }
            if (now != e->slots) {
                std::vector<cohcuda_chan_state> st(e->n_signal + 1);
                if (cohcuda_get_state(e->h, st.data(), e->n_signal + 1) != COHCUDA_OK) die("cohcuda_get_state", e->h);
                j = 1;
                for (auto *d : *ctx->devices) {
                    if (j > e->n_signal) break;
                    st[j].phasecorr[0] = d->phasecorr.real(); st[j].phasecorr[1] = d->phasecorr.imag();
                    st[j].phasecorrprev[0] = d->phasecorrprev.real(); st[j].phasecorrprev[1] = d->phasecorrprev.imag();
                    j++;
                }
                if (cohcuda_set_state(e->h, st.data(), e->n_signal + 1) != COHCUDA_OK) die("cohcuda_set_state", e->h);
                e->slots = now;
            }
        }

        // cbuffer::setbufferptr's in-place XOR (include/common.h:121) is deferred to the kernels' load path
        // (host/cdsp.cc): the buffers are still offset-binary unless COHCUDA_HOST_XOR_ON_CALL=1
        const int deferred = cohcuda_host_xor_deferred();
        int injected = g_inject_failures.load();
        while (injected > 0 && !g_inject_failures.compare_exchange_weak(injected, injected - 1)) {}
        if (injected > 0 ||
            cohcuda_submit_block(e->h, e->chan.data(), e->readcnt.data(), e->request.data(),
                                 ctx->refnoise->isenabled() ? 1 : 0, /*input_is_signed=*/deferred ? 0 : 1) != COHCUDA_OK) {
            drop_block(ctx, e, injected > 0 ? "injected failure (test hook)" : "cohcuda_submit_block");
            continue;                                                      // lag requests stay pending (csdrdevice.h:129-151)
        }
        if (cohcuda_wait_block(e->h, &e->res) != COHCUDA_OK) { drop_block(ctx, e, "cohcuda_wait_block"); continue; }
        e->have_result = true; e->consecutive = 0; e->rebuilt = false;

        // reference column: raw signed bytes (:254-255) -- the library's column 0 when the XOR was deferred
        ctx->refdev->packetize.write(0, e->readcnt[0], deferred ? const_cast<int8_t *>(e->res.packet) : refsptr);
        ctx->refdev->consume();

        int c = 1; k = 1;
        for (auto *d : *ctx->devices) {
            if (e->ready[k]) {
                const std::complex<float> p(e->res.phasecorr[2 * k], e->res.phasecorr[2 * k + 1]);
                d->phasecorr = p;                                          // csdrdevice.cc:67-68 (needs the friend line)
                d->phasecorrprev = p;
                int8_t *col = const_cast<int8_t *>(e->res.packet) + (size_t)k * B;
                d->packetize.write(c, e->readcnt[k], col);                 // :278, already packed to int8
                d->packetize.writedebug(c, p);                             // :279
                c++;
                d->consume();                                              // :281
            }
            k++;
        }
        if (ctx->lagqueuesize() > 1) ctx->computelag();                    // :284-285
        ctx->refdev->packetize.notifysend();                               // :288
    }
}

// src/ccoherent.cc:296-302, :240-242
void ccoherent::start() { thread = std::thread(&ccoherent::threadf, this); }
void ccoherent::request_exit() { do_exit = true; }
void ccoherent::join() { thread.join(); }
