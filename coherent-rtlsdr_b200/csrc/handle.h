// handle.h -- the engine handle behind cohcuda_t and the small helpers every api_*.inc uses.
#pragma once
#include "coherent_cuda.h"
#include "kernels.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace coh;

struct GatherCtx;                                   // api_gather.inc

struct cohcuda {
    cohcuda_config cfg{};
    uint32_t nch = 0, B = 0, L = 0;
    size_t hdr = 0, plen = 0, pitch = 0, lead = 0;
    int sm_count = 0, lookback = 48;
    cudaStream_t stream = nullptr, s_in = nullptr, s_out = nullptr;
    bool own_stream = false;
    std::string err;
    uint64_t launches = 0;

    float2 *d_TN = nullptr, *d_TN_sub = nullptr;   // W_N^n (n < L) and, for L > 8192, W_16384^n (n < 8192)
    float4 *d_scratch = nullptr;                   // L > 8192: [corr_chunk * C][L] packed elements (fft_big.cu)
    uint32_t corr_chunk = 0;                       // blocks per K1/K2 launch group
    uint32_t big_items = 0;                        // L > 8192: items per launch group the scratch holds (0 on the small path)
    uint32_t pipe_chunk_cfg = 0;                   // cohcuda_config.pipe_chunk_blocks (0 = default)
    size_t readcnt_cap = 0;                        // words d_readcnt holds
    ChanState *d_state[2] = { nullptr, nullptr };
    int cur = 0;
    uint32_t *d_ticket = nullptr; uint32_t epoch = 0;

    uint32_t cap = 0;                                   // capacity in blocks
    uint8_t *d_iq = nullptr; int8_t *d_packets = nullptr;
    // second set for host-buffer batches: batch k+1's H2D runs while batch k's packets are still
    // being read back (the two copy engines stay busy across the batch boundary)
    uint8_t *d_iq_alt = nullptr; int8_t *d_packets_alt = nullptr;
    int pp = 0;                                    // set the next host-buffer batch uses
    uint32_t alt_cap = 0;                          // blocks the second set holds
    cudaEvent_t ev_comp[2] = {nullptr, nullptr};   // kernels of the last batch that read d_iq[set] are done
    cudaEvent_t ev_outp[2] = {nullptr, nullptr};   // read-back of the last batch that wrote d_packets[set] is done
    float2 *d_refspec = nullptr; PeakRec *d_peaks = nullptr;
    float *d_lag = nullptr, *d_mag = nullptr; float2 *d_phase = nullptr; uint8_t *d_meas = nullptr;
    uint4 *d_q = nullptr;
    int32_t *d_src = nullptr; uint32_t *d_items = nullptr, *d_reflist = nullptr, *d_slot = nullptr;
    uint8_t *d_refnoise = nullptr; uint32_t *d_readcnt = nullptr;
    const int8_t *last_packets = nullptr;
    const uint8_t *last_iq = nullptr; uint32_t last_nb = 0;   // device input and size of the last batch

    // pinned staging for the per-block streaming calls
    uint8_t *h_flags = nullptr; uint32_t *h_readcnt = nullptr;
    float *h_lag = nullptr, *h_mag = nullptr, *h_phase = nullptr; uint8_t *h_meas = nullptr;
    int8_t *h_packet[2] = { nullptr, nullptr };
    int pk = 0; bool pending = false; uint32_t seq = 0;
    std::vector<cudaEvent_t> ev_in, ev_done;
    cudaEvent_t ev_start = nullptr, ev_out = nullptr;
    std::vector<void *> registered;
    // optional per-kernel timing (cohcuda_profile): 0 = k_ref_spectrum, 1 = k_correlate, 2 = k_apply
    bool prof = false; cudaEvent_t pe0 = nullptr, pe1 = nullptr;
    double prof_ms[3] = { 0, 0, 0 }; uint32_t prof_n[3] = { 0, 0, 0 };
    GatherCtx *gather = nullptr;                   // multi-GPU gather state (cohcuda_gather_init), else NULL
};

static constexpr uint32_t kCorrChunk = 1024;            // blocks per K1/K2 launch (device-resident batches): one spectrum slot of 16 L bytes each;
                                                        // measured 148 -> 1024: +3 % (fewer launch tails), the spectra then live in HBM not L2
static constexpr uint32_t kPipeChunk = 148;             // same, when host copies are pipelined around the kernels (swept 24..512 with
                                                        // batches overlapping: 148 best, 7.99 ms per 1024-block URA21 step)

// blocks per launch group of the large-block path (fft_big.cu) for an engine of n_signal channels: the scratch
// holds big_items items.  A function of n_signal only, so that every rank of a channel-sharded gather can work
// out every other rank's value (api_gather.inc)
static inline uint32_t big_corr_chunk(uint32_t big_items, uint32_t n_signal) {
    const uint32_t cc = big_items / (n_signal ? n_signal : 1);
    return cc < 1 ? 1 : (cc > kCorrChunk ? kCorrChunk : cc);
}

#define CK(call)                                                                        \
    do {                                                                                \
        cudaError_t e__ = (call);                                                       \
        if (e__ != cudaSuccess) {                                                       \
            h->err = std::string(#call) + ": " + cudaGetErrorString(e__);               \
            return COHCUDA_ERR_CUDA;                                                    \
        }                                                                               \
    } while (0)

template <typename T> static cudaError_t dalloc(T **p, size_t n) {
    cudaError_t e = cudaMalloc((void **)p, n * sizeof(T));
    if (e == cudaSuccess) e = cudaMemset(*p, 0, n * sizeof(T));
    return e;
}
template <typename T> static void dfree(T *&p) { if (p) cudaFree(p); p = nullptr; }

