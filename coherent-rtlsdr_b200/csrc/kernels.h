// kernels.h -- internal launch interfaces shared by the .cu files of libcoherent_cuda.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace coh {

struct PeakRec { uint32_t idx; float val; };      // arg-max bin and |c|^2 there (ccoherent.cc:192,203)

struct ChanState {                                // device mirror of cohcuda_chan_state
    float2  phasecorr, phasecorrprev;
    float   lag, mag;
    int32_t held_lag;
    int32_t pad;
};

struct CorrLaunch {
    const uint8_t *iq; uint32_t nch, B;
    uint32_t n_ref_blocks; const uint32_t *ref_block_list;   // blocks whose ref spectrum is needed (NULL = 0..n-1)
    uint32_t n_items; const uint32_t *item_list;             // b*nch+c items (NULL = all signal channels of all blocks)
    const uint32_t *block_slot;                              // block -> refspec slot (NULL = identity)
    uint32_t xormask;
    const float2 *TN; float2 *refspec; PeakRec *peaks;
    // blocks above 8192 samples (fft_big.cu): HBM scratch [n_items][L] float4 and the twiddle
    // table of the 8192-point sub-transforms
    float4 *scratch; const float2 *TN_sub;
    int sm_count;
    uint32_t *ticket;               // device counter of the persistent L = 8192 kernel (NULL: one CTA per item)
};
cudaError_t launch_correlate(const CorrLaunch &p, cudaStream_t st);
cudaError_t launch_correlate_big(const CorrLaunch &p, cudaStream_t st);
bool correlate_supported(uint32_t B);
bool correlate_is_big(uint32_t B);

// sample covariance sums of packet columns (cov_kernel.cu); S [blocks][nch][nch][2] int64, m [blocks][nch][2] int32
cudaError_t launch_covariance(const int8_t *packets, size_t pitch, size_t lead, size_t hdr, uint32_t nch, uint32_t B,
                              uint32_t n_blocks, long long *S, int *m, cudaStream_t st);

// fractional delay by parabolic interpolation of exact integer correlation values (frac_kernel.cu)
cudaError_t launch_fraclag(const uint8_t *iq, uint32_t nch, uint32_t B, uint32_t n_blocks, uint32_t xormask,
                           const float *lag, float *frac, double *y3, cudaStream_t st);

struct ApplyLaunch {
    const uint8_t *iq; uint32_t nch, B, n_blocks;     // n_blocks = whole batch (for the state hand-over)
    uint32_t block0, n_launch_blocks;                  // blocks [block0, block0+n_launch_blocks) in this launch
    uint32_t xormask;               // 0x80808080 for offset-binary input, 0 if already signed
    int apply_shift, noheader;
    float alpha; int lookback;      // IIR weight and look-back depth (enabled blocks)
    // lag source: src_block[b*nch+c] = block whose peak applies (-1 = carried state); NULL with
    // lag_all=1 means "own block", NULL with lag_all=0 means "carried state"
    const int32_t *src_block; int lag_all;
    const PeakRec *peaks;
    const uint8_t *refnoise; int refnoise_scalar;            // per-block enable flags or scalar
    const uint32_t *readcnt; uint32_t readcnt0, globalseqn0;
    const ChanState *state_in; ChanState *state_out;
    int8_t *packets; size_t pitch, lead;                     // device packet layout
    // where this launch's rows land inside the output packets (multi-GPU gather, api_gather.inc).
    // Defaults (out_nch = nch, out_col0 = 1, out_block0 = 0, both flags) = one rank's own packets.
    uint32_t out_nch;               // columns of the output packet: header size 16 + 4*out_nch, hdr0.N
    uint32_t out_col0;              // output column of local signal channel 1
    uint32_t out_block0;            // output packet row of local block 0
    int write_ref, write_hdr;       // this rank writes column 0 / the header + readcnt words
    uint32_t out_skip;              // local blocks < out_skip (a shard's warm-up halo) are not part of the output:
    int8_t *packets_skip;           //   their rows go to this local scratch (row b, same pitch/lead) instead
    uint32_t max_ctas;              // grid cap (0 = one CTA per SM): a k_apply that stores over NVLink runs on a few SMs
    float *lag, *mag; float2 *phase; uint8_t *measured;      // per (block, channel) outputs
    float2 *cf32;                                            // optional corrected float matrix
    // look-back scratch
    uint4 *q; uint32_t epoch;               // 16-byte records {a, b, epoch, check}
    uint32_t *ticket;                       // device counter (0 between launches)
    int sm_count;
};
cudaError_t launch_apply(const ApplyLaunch &p, cudaStream_t st);

// cdsp primitive kernels (device pointers)
cudaError_t prim_convtosigned(const uint8_t *in, uint8_t *out, int n, cudaStream_t st);
cudaError_t prim_convtofloat(float *out, const int8_t *in, int n, cudaStream_t st);
cudaError_t prim_convto8bit(int8_t *out, const float *in, int nfloats, cudaStream_t st);
cudaError_t prim_scalarmul(float2 *out, const float2 *in, float2 s, int n, cudaStream_t st);
cudaError_t prim_conj_dot(float2 *res, const float2 *a, const float2 *b, int n, cudaStream_t st);
cudaError_t prim_magsquared(float *out, const float2 *in, int n, cudaStream_t st);
cudaError_t prim_conjugatemul(float2 *out, const float2 *a, const float2 *b, int n, cudaStream_t st);
cudaError_t prim_indexofmax(uint32_t *idx, const float *in, int n, cudaStream_t st);
cudaError_t prim_sumsq(float *res, const float *in, int n, cudaStream_t st);
cudaError_t prim_fft(float2 *out, const float2 *in, int n, int howmany, int sign, const float2 *TN_n, cudaStream_t st);
bool prim_fft_supported(int n);

}  // namespace coh
