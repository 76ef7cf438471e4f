// lag_static.cu -- cohcuda_lags_static(): the lag measurement (reference spectrum -> correlate ->
// arg-max, src/ccoherent.cc:174-232 in the reference tree) for callers that must not allocate.
//
// Every other entry point works on a cohcuda_t handle, and cohcuda_create()/cohcuda_batch() obtain
// their working memory (twiddle tables, the reference-spectrum slots, the peak records, the ticket
// of the persistent kernel) with cudaMalloc.  Three kinds of users of this library cannot do that:
//   * a receiver's DSP thread that has a block period to keep (8192 samples at 2.4 MS/s = 3.4 ms):
//     cudaMalloc/cudaFree serialise against the whole device and can stall it for longer than that;
//   * a plug-in inside a host program whose own pool has already claimed the device's memory, so a
//     late cudaMalloc fails at an arbitrary block instead of at load time;
//   * code that runs where the CUDA allocator is off limits by policy (a safety-reviewed acquisition
//     loop with a fixed memory map, or a process that only holds memory handed to it by a parent).
// For them the working memory is the fixed array g_lag_scratch below: it is part of the library's
// device image, so every GPU that loads libcoherent_cuda.so reserves kScratchBytes (about 4.2 MiB)
// from the first CUDA call of the process until exit, whether or not this entry point is ever used.
// The price is a size limit, stated in include/coherent_cuda.h and enforced here.
//
// The kernels are the ones cohcuda_batch() launches (launch_correlate, fft_kernels.cu), given pointers
// into the array instead of cudaMalloc'ed buffers; only the last, trivial step differs: a small kernel
// turns the peak records into lag / mag with the arithmetic k_apply uses (apply_kernel.cu, "lag that
// applies to this block"), so that the two entry points agree bit for bit.
#include "coherent_cuda.h"
#include "kernels.h"
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <mutex>
#include <vector>

namespace {

using namespace coh;

// ---- layout of the fixed working memory ------------------------------------------------------------
// [ticket 256 B][twiddle tables of every supported block size][peak records][reference spectra]
constexpr uint32_t kMaxB = COHCUDA_STATIC_MAX_BLOCKSIZE;                  // 16384 bytes = 8192 samples
constexpr uint32_t kMinB = 64;
constexpr size_t kSpecBytes = COHCUDA_STATIC_MAX_SPECTRUM_BYTES;          // 16 L = 8 B bytes per block
constexpr uint32_t kMaxItems = COHCUDA_STATIC_MAX_ITEMS;                  // n_blocks * (n_signal + 1)
constexpr size_t kOffTicket = 0;
constexpr size_t kOffTw = 256;
// table of block size B holds L = B/2 float2 (W_2L^n, n < L); tables of L = 32, 64, ... 8192 back to back:
// the table of L starts (L - 32) entries in, (2 * 8192 - 32) entries altogether
constexpr size_t kTwEntries = 2 * (kMaxB / 2) - 32;
constexpr size_t kOffPeaks = (kOffTw + kTwEntries * sizeof(float2) + 255) / 256 * 256;
constexpr size_t kOffSpec = (kOffPeaks + (size_t)kMaxItems * sizeof(PeakRec) + 255) / 256 * 256;
constexpr size_t kScratchBytes = kOffSpec + kSpecBytes;

// In host code this name is only a handle for cudaGetSymbolAddress: it is NOT a device address and must
// never be handed to a kernel or a copy directly.  device_setup() below resolves it per device.
static __device__ __align__(256) unsigned char g_lag_scratch[kScratchBytes];

struct DevSetup {
    bool ready = false;
    unsigned char *base = nullptr;       // device address of g_lag_scratch on this device
    int sm_count = 0;
};
DevSetup g_setup[64];
std::mutex g_setup_mutex;

int fail(char *errbuf, size_t errbuf_len, int status, const char *fmt, ...) __attribute__((format(printf, 4, 5)));
int fail(char *errbuf, size_t errbuf_len, int status, const char *fmt, ...) {
    if (errbuf && errbuf_len) {
        va_list ap; va_start(ap, fmt);
        vsnprintf(errbuf, errbuf_len, fmt, ap);
        va_end(ap);
    }
    return status;
}

// Once per device, with that device current: resolve the array's device address and fill the twiddle
// tables (the same values cohcuda_create computes: cos/sin in double, rounded once).  Synchronous --
// this is why the first call on a device belongs outside any stream capture.
cudaError_t device_setup(int dev, DevSetup **out) {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    DevSetup &s = g_setup[dev];
    *out = &s;
    if (s.ready) return cudaSuccess;
    unsigned char *base = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&base, g_lag_scratch);
    if (e != cudaSuccess) return e;
    if (!base) return cudaErrorInvalidSymbol;
    std::vector<float2> tw(kTwEntries);                      // host memory only
    for (uint32_t L = kMinB / 2; L <= kMaxB / 2; L <<= 1)
        for (uint32_t n = 0; n < L; n++) {
            const double a = -2.0 * M_PI * (double)n / (double)(2 * L);
            tw[(L - 32) + n] = make_float2((float)std::cos(a), (float)std::sin(a));
        }
    e = cudaMemcpy(base + kOffTw, tw.data(), kTwEntries * sizeof(float2), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return e;
    e = cudaDeviceSynchronize();                             // the pageable copy has landed before any stream uses it
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&s.sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    s.base = base;
    s.ready = true;
    return cudaSuccess;
}

bool on_device(const void *p) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

// peak record -> what set_lag() publishes (include/csdrdevice.h:138-151): the same two expressions as
// k_apply (ccoherent.cc:203-204, 218-232); the reference column reports 0 / 0 as cohcuda_batch does
__global__ void k_peaks_to_lags(const PeakRec *__restrict__ peaks, uint32_t nch, uint32_t L, uint32_t n,
                                float *__restrict__ lag, float *__restrict__ mag) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float lagf = 0.0f, magf = 0.0f;
    if (i % nch != 0) {
        const PeakRec pk = peaks[i];
        lagf = (float)pk.idx - (float)L;
        magf = __fsqrt_rn(__fdiv_rn(pk.val, (float)L));
    }
    lag[i] = lagf;
    mag[i] = magf;
}

}  // namespace

// See include/coherent_cuda.h for the contract.  No cudaMalloc / cudaFree / cudaMallocAsync and no
// library that allocates is reached from here: launch_correlate() only sets function attributes, clears
// the 4-byte ticket with cudaMemsetAsync and launches.
extern "C" int cohcuda_lags_static(const uint8_t *iq_dev, uint32_t n_signal, uint32_t blocksize, uint32_t n_blocks,
                                   int input_is_signed, float *lag_dev, float *mag_dev, void *stream,
                                   char *errbuf, size_t errbuf_len) {
    if (errbuf && errbuf_len) errbuf[0] = 0;
    if (!iq_dev || !lag_dev || !mag_dev)
        return fail(errbuf, errbuf_len, COHCUDA_ERR_ARG, "cohcuda_lags_static: iq_dev, lag_dev and mag_dev must be non-null device pointers");
    if (blocksize < kMinB || (blocksize & (blocksize - 1)))
        return fail(errbuf, errbuf_len, COHCUDA_ERR_ARG, "cohcuda_lags_static: blocksize %u is not a power of two >= %u", blocksize, kMinB);
    if (blocksize > kMaxB)
        return fail(errbuf, errbuf_len, COHCUDA_ERR_UNSUPPORTED,
                    "cohcuda_lags_static: blocksize %u exceeds COHCUDA_STATIC_MAX_BLOCKSIZE = %u (larger blocks need per-item "
                    "scratch that does not fit the fixed array: use cohcuda_batch)", blocksize, kMaxB);
    if (n_signal == 0 || n_blocks == 0)
        return fail(errbuf, errbuf_len, COHCUDA_ERR_ARG, "cohcuda_lags_static: n_signal and n_blocks must be at least 1");
    const uint64_t items = (uint64_t)n_blocks * ((uint64_t)n_signal + 1);
    if (items > kMaxItems)
        return fail(errbuf, errbuf_len, COHCUDA_ERR_UNSUPPORTED,
                    "cohcuda_lags_static: n_blocks * (n_signal + 1) = %llu exceeds COHCUDA_STATIC_MAX_ITEMS = %u: split the call",
                    (unsigned long long)items, kMaxItems);
    const uint64_t spec = (uint64_t)n_blocks * 8ull * blocksize;
    if (spec > kSpecBytes)
        return fail(errbuf, errbuf_len, COHCUDA_ERR_UNSUPPORTED,
                    "cohcuda_lags_static: n_blocks * 8 * blocksize = %llu bytes of reference spectra exceed "
                    "COHCUDA_STATIC_MAX_SPECTRUM_BYTES = %llu (at most %llu blocks of %u bytes per call)",
                    (unsigned long long)spec, (unsigned long long)kSpecBytes, (unsigned long long)(kSpecBytes / (8ull * blocksize)), blocksize);
    if ((uintptr_t)iq_dev & 15)
        return fail(errbuf, errbuf_len, COHCUDA_ERR_ARG, "cohcuda_lags_static: iq_dev must be 16-byte aligned");

    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return fail(errbuf, errbuf_len, COHCUDA_ERR_CUDA, "cohcuda_lags_static: no usable CUDA device (there is no CPU fallback): %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= 64)
        return fail(errbuf, errbuf_len, COHCUDA_ERR_UNSUPPORTED, "cohcuda_lags_static: device ordinal %d is outside 0..63", dev);
    if (!on_device(iq_dev) || !on_device(lag_dev) || !on_device(mag_dev))
        return fail(errbuf, errbuf_len, COHCUDA_ERR_ARG, "cohcuda_lags_static: iq_dev, lag_dev and mag_dev must point to device memory");
    DevSetup *s = nullptr;
    e = device_setup(dev, &s);
    if (e != cudaSuccess)
        return fail(errbuf, errbuf_len, COHCUDA_ERR_CUDA, "cohcuda_lags_static: set-up of the fixed working memory on device %d: %s", dev, cudaGetErrorString(e));

    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t nch = n_signal + 1, L = blocksize / 2;
    CorrLaunch cl{};
    cl.iq = iq_dev; cl.nch = nch; cl.B = blocksize;
    cl.xormask = input_is_signed ? 0u : 0x8080u;
    cl.TN = reinterpret_cast<const float2 *>(s->base + kOffTw) + (L - 32);
    cl.refspec = reinterpret_cast<float2 *>(s->base + kOffSpec);
    cl.peaks = reinterpret_cast<PeakRec *>(s->base + kOffPeaks);
    cl.ticket = reinterpret_cast<uint32_t *>(s->base + kOffTicket);      // zeroed by launch_correlate before each use
    cl.sm_count = s->sm_count;
    cl.n_ref_blocks = n_blocks; cl.n_items = n_blocks * n_signal;
    (void)cudaGetLastError();                                            // a stale non-sticky error is not ours to report
    e = launch_correlate(cl, st);
    if (e != cudaSuccess)
        return fail(errbuf, errbuf_len, COHCUDA_ERR_CUDA, "cohcuda_lags_static: correlation launch: %s", cudaGetErrorString(e));
    const uint32_t n = (uint32_t)items;
    k_peaks_to_lags<<<(n + 255) / 256, 256, 0, st>>>(cl.peaks, nch, L, n, lag_dev, mag_dev);
    e = cudaGetLastError();
    if (e != cudaSuccess)
        return fail(errbuf, errbuf_len, COHCUDA_ERR_CUDA, "cohcuda_lags_static: k_peaks_to_lags launch: %s", cudaGetErrorString(e));
    return COHCUDA_OK;
}
