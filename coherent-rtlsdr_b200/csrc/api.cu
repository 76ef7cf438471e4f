// api.cu -- C ABI of libcoherent_cuda (include/coherent_cuda.h): handle, buffers, the batch
// schedule (K1 ref spectrum -> K2 correlate -> K3 apply) and the per-block streaming calls.
// Host-side only; kernels live in fft_kernels.cu / apply_kernel.cu / prims.cu.
#include "handle.h"

static std::string g_create_error;

#include "api_gather.inc"

static void free_batch_buffers(cohcuda *h) {
    dfree(h->d_iq); dfree(h->d_packets); dfree(h->d_iq_alt); dfree(h->d_packets_alt); dfree(h->d_refspec); dfree(h->d_peaks); dfree(h->d_scratch);
    dfree(h->d_lag); dfree(h->d_mag); dfree(h->d_phase); dfree(h->d_meas);
    dfree(h->d_q);
    dfree(h->d_src); dfree(h->d_items); dfree(h->d_reflist); dfree(h->d_slot);
    dfree(h->d_refnoise); dfree(h->d_readcnt);
    h->cap = 0;
}

static int ensure_capacity(cohcuda *h, uint32_t nb) {
    if (nb <= h->cap) return COHCUDA_OK;
    CK(cudaStreamSynchronize(h->s_in));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaStreamSynchronize(h->s_out));
    free_batch_buffers(h);
    const size_t n = (size_t)nb * h->nch;
    CK(dalloc(&h->d_iq, n * h->B));
    CK(dalloc(&h->d_packets, (size_t)nb * h->pitch + 16));
    h->alt_cap = 0;                                  // the second set is allocated by the first host-buffer batch
    if (correlate_supported(h->B)) {
        const uint32_t cc = h->corr_chunk < nb ? h->corr_chunk : nb;       // no more slots than blocks
        CK(dalloc(&h->d_refspec, (size_t)cc * 2 * h->L));
        if (correlate_is_big(h->B)) CK(dalloc(&h->d_scratch, (size_t)cc * h->cfg.n_signal * h->L));
    }
    CK(dalloc(&h->d_peaks, n));
    CK(dalloc(&h->d_lag, n)); CK(dalloc(&h->d_mag, n)); CK(dalloc(&h->d_phase, n)); CK(dalloc(&h->d_meas, n));
    CK(dalloc(&h->d_q, n));
    CK(dalloc(&h->d_src, n)); CK(dalloc(&h->d_items, n)); CK(dalloc(&h->d_reflist, (size_t)nb)); CK(dalloc(&h->d_slot, (size_t)nb));
    CK(dalloc(&h->d_refnoise, (size_t)nb)); CK(dalloc(&h->d_readcnt, n)); h->readcnt_cap = n;
    h->cap = nb;
    return COHCUDA_OK;
}

extern "C" const char *cohcuda_last_error(cohcuda_t *h) { return h ? h->err.c_str() : g_create_error.c_str(); }

extern "C" int cohcuda_create(const cohcuda_config *cfg, cohcuda_t **out) {
    if (!cfg || !out) { g_create_error = "null argument"; return COHCUDA_ERR_ARG; }
    *out = nullptr;
    const uint32_t B = cfg->blocksize;
    if (B < 64 || (B & (B - 1)) || cfg->n_signal == 0 || cfg->n_signal > 65535) {
        g_create_error = "blocksize must be a power of two >= 64 and n_signal in 1..65535";
        return COHCUDA_ERR_ARG;
    }
    // the phase IIR is evaluated by a look-back of ceil(48 ln2 / -ln(1-alpha)) blocks, at most 2048:
    // below alpha = 1 - 2^(-48/2048) the truncated tail would no longer be below fp32 resolution
    if (!(cfg->alpha >= 0.0162f && cfg->alpha <= 1.0f)) { g_create_error = "alpha must be in [0.0162, 1] (look-back depth <= 2048 blocks)"; return COHCUDA_ERR_ARG; }
    if (B > (1u << 21)) { g_create_error = "blocksize above 2 MiB is not supported (32-bit lanes of the exact phase dot product)"; return COHCUDA_ERR_ARG; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || cfg->device < 0 || cfg->device >= ndev) {
        g_create_error = std::string("no usable CUDA device (there is no CPU fallback): ") +
                         (e != cudaSuccess ? cudaGetErrorString(e) : "device ordinal out of range");
        return COHCUDA_ERR_CUDA;
    }
    cohcuda *h = new (std::nothrow) cohcuda();
    if (!h) { g_create_error = "out of host memory"; return COHCUDA_ERR_NOMEM; }
    h->cfg = *cfg; h->B = B; h->L = B / 2; h->nch = cfg->n_signal + 1;
    h->hdr = cfg->noheader ? 0 : 16 + 4 * (size_t)h->nch;
    h->plen = h->hdr + (size_t)h->nch * B;
    h->lead = (16 - h->hdr % 16) % 16;
    h->pitch = (h->lead + h->plen + 15) / 16 * 16;
    // look-back depth: truncated IIR tail weight (1-alpha)^depth <= 2^-48
    if (cfg->alpha >= 1.0f) h->lookback = 1;
    else {
        double d = std::ceil(48.0 * std::log(2.0) / -std::log(1.0 - (double)cfg->alpha));
        h->lookback = (int)std::fmin(std::fmax(d, 1.0), 2048.0);
    }
    auto fail = [&](const char *what, cudaError_t ce) {
        g_create_error = std::string(what) + ": " + cudaGetErrorString(ce);
        cohcuda_destroy(h);
        return COHCUDA_ERR_CUDA;
    };
    if ((e = cudaSetDevice(cfg->device)) != cudaSuccess) return fail("cudaSetDevice", e);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, cfg->device)) != cudaSuccess) return fail("cudaGetDeviceProperties", e);
    h->sm_count = prop.multiProcessorCount;
    if (cfg->stream) h->stream = (cudaStream_t)cfg->stream;
    else {
        if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
        h->own_stream = true;
    }
    if ((e = cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    if ((e = cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    if ((e = cudaEventCreateWithFlags(&h->ev_start, cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaEventCreateWithFlags(&h->ev_out, cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
    for (int i = 0; i < 2; i++) {
        if ((e = cudaEventCreateWithFlags(&h->ev_comp[i], cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
        if ((e = cudaEventCreateWithFlags(&h->ev_outp[i], cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
    }
    // twiddle table: exp(-2*pi*i*n/N), n < L, N = 2L, computed in double and rounded once
    {
        std::vector<float2> tn(h->L);
        for (uint32_t n = 0; n < h->L; n++) {
            double a = -2.0 * M_PI * (double)n / (double)(2 * h->L);
            tn[n] = make_float2((float)std::cos(a), (float)std::sin(a));
        }
        if ((e = cudaMalloc((void **)&h->d_TN, sizeof(float2) * h->L)) != cudaSuccess) return fail("cudaMalloc", e);
        if ((e = cudaMemcpy(h->d_TN, tn.data(), sizeof(float2) * h->L, cudaMemcpyHostToDevice)) != cudaSuccess) return fail("cudaMemcpy", e);
    }
    h->corr_chunk = kCorrChunk;
    if (cfg->pipe_chunk_blocks > kCorrChunk) { g_create_error = "pipe_chunk_blocks must be at most 1024"; cohcuda_destroy(h); return COHCUDA_ERR_ARG; }
    h->pipe_chunk_cfg = cfg->pipe_chunk_blocks;
    if (correlate_supported(B) && correlate_is_big(B)) {
        // the split path keeps 16 L bytes of scratch per item in flight: bound the items per launch group
        const uint32_t big_items = 2048;               // swept 256..5376 at L = 65536: 45.6 -> 52.2 G samples/s (2 GiB of scratch)
        h->big_items = big_items;
        h->corr_chunk = big_corr_chunk(big_items, cfg->n_signal);
        std::vector<float2> ts(8192);
        for (uint32_t n = 0; n < 8192; n++) {
            double a = -2.0 * M_PI * (double)n / 16384.0;
            ts[n] = make_float2((float)std::cos(a), (float)std::sin(a));
        }
        if ((e = cudaMalloc((void **)&h->d_TN_sub, sizeof(float2) * 8192)) != cudaSuccess) return fail("cudaMalloc", e);
        if ((e = cudaMemcpy(h->d_TN_sub, ts.data(), sizeof(float2) * 8192, cudaMemcpyHostToDevice)) != cudaSuccess) return fail("cudaMemcpy", e);
    }
    for (int i = 0; i < 2; i++)
        if ((e = dalloc(&h->d_state[i], (size_t)h->nch)) != cudaSuccess) return fail("cudaMalloc", e);
    if ((e = dalloc(&h->d_ticket, (size_t)4)) != cudaSuccess) return fail("cudaMalloc", e);   // [0] k_apply, [1] k_correlate
    {
        std::vector<ChanState> st(h->nch);
        for (auto &s : st) {            // csdrdevice.cc:36-40
            s.phasecorr = make_float2(1.f, 0.f); s.phasecorrprev = make_float2(1.f, 0.f);
            s.lag = 0.f; s.mag = 0.f; s.held_lag = 0; s.pad = 0;
        }
        for (int i = 0; i < 2; i++)
            if ((e = cudaMemcpy(h->d_state[i], st.data(), sizeof(ChanState) * h->nch, cudaMemcpyHostToDevice)) != cudaSuccess) return fail("cudaMemcpy", e);
    }
    const size_t n1 = h->nch;
    if ((e = cudaMallocHost((void **)&h->h_flags, n1)) != cudaSuccess) return fail("cudaMallocHost", e);
    if ((e = cudaMallocHost((void **)&h->h_readcnt, 4 * n1)) != cudaSuccess) return fail("cudaMallocHost", e);
    if ((e = cudaMallocHost((void **)&h->h_lag, 4 * n1)) != cudaSuccess) return fail("cudaMallocHost", e);
    if ((e = cudaMallocHost((void **)&h->h_mag, 4 * n1)) != cudaSuccess) return fail("cudaMallocHost", e);
    if ((e = cudaMallocHost((void **)&h->h_phase, 8 * n1)) != cudaSuccess) return fail("cudaMallocHost", e);
    if ((e = cudaMallocHost((void **)&h->h_meas, n1)) != cudaSuccess) return fail("cudaMallocHost", e);
    for (int i = 0; i < 2; i++)
        if ((e = cudaMallocHost((void **)&h->h_packet[i], h->plen)) != cudaSuccess) return fail("cudaMallocHost", e);
    int rc = ensure_capacity(h, cfg->max_batch_blocks ? cfg->max_batch_blocks : 1);
    if (rc != COHCUDA_OK) { g_create_error = h->err; cohcuda_destroy(h); return rc; }
    *out = h;
    return COHCUDA_OK;
}

extern "C" void cohcuda_destroy(cohcuda_t *h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    cohcuda_gather_destroy(h);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (void *p : h->registered) cudaHostUnregister(p);
    free_batch_buffers(h);
    dfree(h->d_TN); dfree(h->d_TN_sub); dfree(h->d_state[0]); dfree(h->d_state[1]); dfree(h->d_ticket);
    if (h->h_flags) cudaFreeHost(h->h_flags);
    if (h->h_readcnt) cudaFreeHost(h->h_readcnt);
    if (h->h_lag) cudaFreeHost(h->h_lag);
    if (h->h_mag) cudaFreeHost(h->h_mag);
    if (h->h_phase) cudaFreeHost(h->h_phase);
    if (h->h_meas) cudaFreeHost(h->h_meas);
    for (int i = 0; i < 2; i++) if (h->h_packet[i]) cudaFreeHost(h->h_packet[i]);
    for (auto ev : h->ev_in) cudaEventDestroy(ev);
    for (auto ev : h->ev_done) cudaEventDestroy(ev);
    if (h->ev_start) cudaEventDestroy(h->ev_start);
    if (h->ev_out) cudaEventDestroy(h->ev_out);
    for (int i = 0; i < 2; i++) { if (h->ev_comp[i]) cudaEventDestroy(h->ev_comp[i]); if (h->ev_outp[i]) cudaEventDestroy(h->ev_outp[i]); }
    if (h->pe0) cudaEventDestroy(h->pe0);
    if (h->pe1) cudaEventDestroy(h->pe1);
    if (h->s_in) cudaStreamDestroy(h->s_in);
    if (h->s_out) cudaStreamDestroy(h->s_out);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

extern "C" int cohcuda_register_host(cohcuda_t *h, void *p, size_t bytes) {
    if (!h || !p || !bytes) return COHCUDA_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
    if (e != cudaSuccess) {
        // pinning is an optimisation (an unpinned buffer is copied synchronously): report the failure and
        // clear the runtime's last-error so that the next kernel launch does not inherit it
        (void)cudaGetLastError();
        h->err = std::string("cudaHostRegister: ") + cudaGetErrorString(e);
        return COHCUDA_ERR_CUDA;
    }
    h->registered.push_back(p);
    return COHCUDA_OK;
}

extern "C" int cohcuda_unregister_host(cohcuda_t *h, void *p) {
    if (!h || !p) return COHCUDA_ERR_ARG;
    for (size_t i = 0; i < h->registered.size(); i++)
        if (h->registered[i] == p) {
            CK(cudaStreamSynchronize(h->stream));
            CK(cudaHostUnregister(p));
            h->registered.erase(h->registered.begin() + i);
            return COHCUDA_OK;
        }
    h->err = "pointer was not registered";
    return COHCUDA_ERR_ARG;
}

extern "C" int cohcuda_packet_layout(cohcuda_t *h, size_t *packet_len, size_t *dev_pitch, size_t *dev_lead) {
    if (!h) return COHCUDA_ERR_ARG;
    if (packet_len) *packet_len = h->plen;
    if (dev_pitch) *dev_pitch = h->pitch;
    if (dev_lead) *dev_lead = h->lead;
    return COHCUDA_OK;
}

extern "C" int cohcuda_sync(cohcuda_t *h) {
    if (!h) return COHCUDA_ERR_ARG;
    CK(cudaStreamSynchronize(h->s_out));
    if (h->gather) CK(cudaStreamSynchronize(h->gather->s_g));
    CK(cudaStreamSynchronize(h->stream));
    return COHCUDA_OK;
}

// stream-ordered join: work enqueued on the handle's stream after this call runs after every
// outstanding packet read-back (no host wait)
extern "C" int cohcuda_flush(cohcuda_t *h) {
    if (!h) return COHCUDA_ERR_ARG;
    CK(cudaEventRecord(h->ev_out, h->s_out));
    CK(cudaStreamWaitEvent(h->stream, h->ev_out, 0));
    if (h->gather) {                                  // and after every transfer of the gather
        CK(cudaEventRecord(h->gather->ev_join, h->gather->s_g));
        CK(cudaStreamWaitEvent(h->stream, h->gather->ev_join, 0));
    }
    return COHCUDA_OK;
}

extern "C" int cohcuda_covariance(cohcuda_t *h, const int8_t *packets_dev, uint32_t n_blocks,
                                  int64_t *S_host, int32_t *m_host) {
    if (!h || !S_host || !m_host) { if (h) h->err = "null argument"; return COHCUDA_ERR_ARG; }
    if (!packets_dev) packets_dev = h->last_packets;
    if (!packets_dev) { h->err = "no packets: run a batch first or pass device packets"; return COHCUDA_ERR_STATE; }
    if (n_blocks == 0) return COHCUDA_OK;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->s_out));
    const size_t nS = (size_t)n_blocks * h->nch * h->nch * 2, nm = (size_t)n_blocks * h->nch * 2;
    long long *dS = nullptr; int *dm = nullptr;
    CK(cudaMalloc((void **)&dS, nS * sizeof(long long)));
    cudaError_t e = cudaMalloc((void **)&dm, nm * sizeof(int));
    if (e == cudaSuccess) e = launch_covariance(packets_dev, h->pitch, h->lead, h->hdr, h->nch, h->B, n_blocks, dS, dm, h->stream);
    if (e == cudaSuccess) { h->launches++; e = cudaMemcpyAsync(S_host, dS, nS * sizeof(long long), cudaMemcpyDeviceToHost, h->stream); }
    if (e == cudaSuccess) e = cudaMemcpyAsync(m_host, dm, nm * sizeof(int), cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(dS); cudaFree(dm);
    CK(e);
    return COHCUDA_OK;
}

extern "C" int cohcuda_fraclag(cohcuda_t *h, const uint8_t *iq_dev, uint32_t n_blocks, int input_is_signed,
                               float *frac_host, double *y3_host) {
    if (!h || !frac_host) { if (h) h->err = "null argument"; return COHCUDA_ERR_ARG; }
    if (!iq_dev) iq_dev = h->last_iq;
    if (!iq_dev || n_blocks > h->last_nb) { h->err = "fraclag: no batch of that size has been run"; return COHCUDA_ERR_STATE; }
    if (n_blocks == 0) return COHCUDA_OK;
    CK(cudaSetDevice(h->cfg.device));
    const size_t n = (size_t)n_blocks * h->nch;
    float *dF = nullptr; double *dY = nullptr;
    CK(cudaMalloc((void **)&dF, n * sizeof(float)));
    cudaError_t e = y3_host ? cudaMalloc((void **)&dY, n * 3 * sizeof(double)) : cudaSuccess;
    if (e == cudaSuccess) e = launch_fraclag(iq_dev, h->nch, h->B, n_blocks, input_is_signed ? 0u : 0x8080u, h->d_lag, dF, dY, h->stream);
    if (e == cudaSuccess) { h->launches++; e = cudaMemcpyAsync(frac_host, dF, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream); }
    if (e == cudaSuccess && y3_host) e = cudaMemcpyAsync(y3_host, dY, n * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(dF); cudaFree(dY);
    CK(e);
    return COHCUDA_OK;
}

extern "C" uint64_t cohcuda_launch_count(cohcuda_t *h) { return h ? h->launches : 0; }
extern "C" const int8_t *cohcuda_device_packets(cohcuda_t *h) { return h ? h->last_packets : nullptr; }

extern "C" int cohcuda_get_state(cohcuda_t *h, cohcuda_chan_state *st, uint32_t n) {
    if (!h || !st || n > h->nch) return COHCUDA_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    std::vector<ChanState> tmp(n);
    CK(cudaMemcpy(tmp.data(), h->d_state[h->cur], sizeof(ChanState) * n, cudaMemcpyDeviceToHost));
    for (uint32_t i = 0; i < n; i++) {
        st[i].phasecorr[0] = tmp[i].phasecorr.x; st[i].phasecorr[1] = tmp[i].phasecorr.y;
        st[i].phasecorrprev[0] = tmp[i].phasecorrprev.x; st[i].phasecorrprev[1] = tmp[i].phasecorrprev.y;
        st[i].lag = tmp[i].lag; st[i].mag = tmp[i].mag; st[i].held_lag = tmp[i].held_lag;
    }
    return COHCUDA_OK;
}

extern "C" int cohcuda_set_state(cohcuda_t *h, const cohcuda_chan_state *st, uint32_t n) {
    if (!h || !st || n > h->nch) return COHCUDA_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    std::vector<ChanState> tmp(n);
    for (uint32_t i = 0; i < n; i++) {
        tmp[i].phasecorr = make_float2(st[i].phasecorr[0], st[i].phasecorr[1]);
        tmp[i].phasecorrprev = make_float2(st[i].phasecorrprev[0], st[i].phasecorrprev[1]);
        tmp[i].lag = st[i].lag; tmp[i].mag = st[i].mag; tmp[i].held_lag = st[i].held_lag; tmp[i].pad = 0;
    }
    CK(cudaMemcpy(h->d_state[h->cur], tmp.data(), sizeof(ChanState) * n, cudaMemcpyHostToDevice));
    return COHCUDA_OK;
}

// run one launch, optionally bracketed by events on the launching stream (profile mode syncs)
template <typename F> static int timed_launch(cohcuda *h, int which, F &&f, cudaStream_t st = nullptr) {
    if (!h->prof) { CK(f()); h->launches++; return COHCUDA_OK; }
    if (!st) st = h->stream;
    if (!h->pe0) { CK(cudaEventCreate(&h->pe0)); CK(cudaEventCreate(&h->pe1)); }
    CK(cudaEventRecord(h->pe0, st));
    CK(f());
    CK(cudaEventRecord(h->pe1, st));
    CK(cudaEventSynchronize(h->pe1));
    float ms = 0; CK(cudaEventElapsedTime(&ms, h->pe0, h->pe1));
    h->prof_ms[which] += ms; h->prof_n[which]++; h->launches++;
    return COHCUDA_OK;
}

extern "C" int cohcuda_profile(cohcuda_t *h, int enable) {
    if (!h) return COHCUDA_ERR_ARG;
    h->prof = enable != 0;
    for (int i = 0; i < 3; i++) { h->prof_ms[i] = 0; h->prof_n[i] = 0; }
    return COHCUDA_OK;
}
extern "C" int cohcuda_profile_read(cohcuda_t *h, double *ms3, uint32_t *launches3) {
    if (!h || !ms3 || !launches3) return COHCUDA_ERR_ARG;
    for (int i = 0; i < 3; i++) { ms3[i] = h->prof_ms[i]; launches3[i] = h->prof_n[i]; }
    return COHCUDA_OK;
}

#include "api_batch.inc"
#include "api_prims.inc"
