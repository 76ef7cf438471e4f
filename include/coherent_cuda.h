/*
 * coherent_cuda.h -- C ABI of libcoherent_cuda.so: the block synchronisation + correction
 * path of coherent-rtlsdr as hand-written sm_100a CUDA.
 *
 * The reference has no FFI: its seam is the C++ class surface of cdsp / ccoherent /
 * cpacketize / csdrdevice.  Each entry point below names the reference code it replaces
 * (file:line in the reference tree); the drop-in C++ classes in coherent-rtlsdr_b200/host/
 * and INTEGRATION.md show the call sites a maintainer would change.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a
 * negative cohcuda_status; nothing throws across the boundary (the reference's DSP path uses
 * int returns and no exceptions, src/crtlsdr.cc:112-135); messages via cohcuda_last_error().
 * One producer thread per handle (the reference has exactly one DSP thread,
 * src/ccoherent.cc:296-298); handles are independent (one per GPU rank).
 * There is no CPU fallback: without a usable CUDA device cohcuda_create() fails.
 */
#ifndef COHERENT_CUDA_H
#define COHERENT_CUDA_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cohcuda cohcuda_t;

typedef enum {
    COHCUDA_OK = 0,
    COHCUDA_ERR_ARG = -1,        /* bad argument / unsupported size          */
    COHCUDA_ERR_CUDA = -2,       /* CUDA runtime error (see last_error)      */
    COHCUDA_ERR_NOMEM = -3,
    COHCUDA_ERR_STATE = -4,      /* wait without submit, etc.                */
    COHCUDA_ERR_UNSUPPORTED = -5
} cohcuda_status;

/* lag_mode for cohcuda_batch / meaning of lag_request == NULL */
#define COHCUDA_LAG_NONE 0   /* TRACK: nobody requests a lag (reference steady state)        */
#define COHCUDA_LAG_ALL  1   /* SYNC: every signal channel requests a lag in every block     */

typedef struct {
    uint32_t n_signal;          /* C, signal channels; the reference channel is extra         */
    uint32_t blocksize;         /* B bytes per channel-block = 2L, power of two, 64..131072 for
                                   the correlation path (src/main.cc:167 default 1<<14)       */
    uint32_t max_lag_slots;     /* reference: nfft-1 = 7 (src/ccoherent.cc:124); 0 = no cap   */
    float    alpha;             /* phase IIR weight; reference 0.5 (src/csdrdevice.cc:60).  Range
                                   [0.0162, 1]: the recurrence is evaluated as a look-back of
                                   ceil(48 ln2 / -ln(1-alpha)) <= 2048 blocks (tail weight 2^-48),
                                   which IS the sequential fp32 result; smaller alpha is rejected */
    int      noheader;          /* -R raw packets (src/cpacketizer.cc:91-96)                  */
    int      apply_shift;       /* 1: software integer-delay removal; 0: as ref   */
    int      device;            /* CUDA device ordinal                                        */
    uint32_t max_batch_blocks;  /* capacity hint for cohcuda_batch (grows on demand)          */
    void    *stream;            /* cudaStream_t to enqueue on, or NULL for a private stream   */
    uint32_t pipe_chunk_blocks; /* tuning: blocks per pipeline stage of host-buffer batches (H2D copy of
                                   stage k+1 | kernels of stage k | read-back of stage k-1); 0 = default
                                   (148 blocks of 22 x 16 KiB, scaled by bytes for other shapes), max 1024.
                                   Results do not depend on it.  The library reads no environment variables. */
} cohcuda_config;

/* per-channel state that survives across blocks (include/csdrdevice.h:74-89) */
typedef struct {
    float   phasecorr[2];       /* csdrdevice::phasecorr                                      */
    float   phasecorrprev[2];   /* csdrdevice::phasecorrprev                                  */
    float   lag;                /* lagpoint.lag (include/csdrdevice.h:42-54)                  */
    float   mag;                /* lagpoint.mag                                               */
    int32_t held_lag;           /* integer lag applied while no new measurement arrives       */
} cohcuda_chan_state;

/* replaces ccoherent::ccoherent buffer/plan set-up (src/ccoherent.cc:32-95) and
 * csdrdevice::csdrdevice state init (src/csdrdevice.cc:20-48) */
int  cohcuda_create(const cohcuda_config *cfg, cohcuda_t **out);
/* replaces ccoherent::~ccoherent (src/ccoherent.cc:97-113) */
void cohcuda_destroy(cohcuda_t *h);
const char *cohcuda_last_error(cohcuda_t *h);   /* h may be NULL: last create() failure */

/* pins librtlsdr-owned USB buffers so block copies are asynchronous; the zero-copy ring of
 * include/common.h:114-122 (USELIBRTLSDRBUFS) hands exactly these pointers to the DSP thread */
int cohcuda_register_host(cohcuda_t *h, void *p, size_t bytes);
int cohcuda_unregister_host(cohcuda_t *h, void *p);

/* packet geometry: packet_len = bytes of one packet as written by cpacketize::write
 * (src/cpacketizer.cc:137-172): [hdr0 16][readcnt 4(C+1)][(C+1) columns x B] (hdr omitted with
 * noheader).  On the device, packet b of a batch starts at base + b*dev_pitch + dev_lead so
 * that every column is 16-byte aligned; host-side packets are contiguous (pitch packet_len). */
int cohcuda_packet_layout(cohcuda_t *h, size_t *packet_len, size_t *dev_pitch, size_t *dev_lead);

typedef struct {
    const float  *lag;          /* C+1 floats (entry 0 = 0): lagpoint.lag per channel         */
    const float  *mag;          /* C+1 floats: lagpoint.mag                                   */
    const float  *phasecorr;    /* 2(C+1) floats: csdrdevice::get_phasecorrect(), what
                                   cpacketize::writedebug publishes (cpacketizer.cc:131-134)  */
    const uint8_t *measured;    /* C+1 flags: 1 where set_lag() would have fired this block   */
    const int8_t *packet;       /* library-owned pinned host memory, valid until the
                                   second-next submit (mirrors packetbuf0/1)                  */
    size_t        packet_len;
    uint32_t      globalseqn;
} cohcuda_result;

/*
 * One block through the schedule of ccoherent::threadf (src/ccoherent.cc:245-294): convert
 * (crtlsdr.cc:205-218), queue + correlate the channels whose lag_request flag is set
 * (ccoherent.cc:123-152,154-239), estimate/apply phase (csdrdevice.cc:58-84), pack into the
 * packet (cpacketizer.cc:137-172).  chan[0] = reference, chan[1..C] = signal; each B bytes,
 * caller-owned, must stay valid until cohcuda_wait_block returns (the reference consume()s
 * right after use, ccoherent.cc:254,281).  input_is_signed = 1 if the in-place XOR of
 * cbuffer::setbufferptr (include/common.h:121) already ran on the buffers.
 * Asynchronous: returns after enqueueing.
 */
int cohcuda_submit_block(cohcuda_t *h, const uint8_t *const *chan, const uint32_t *readcnt,
                         const uint8_t *lag_request, int refnoise_enabled, int input_is_signed);
/* blocks until the submitted block is done; replaces the tail of threadf up to notifysend()
 * (ccoherent.cc:284-288) and fills what set_lag() publishes (csdrdevice.h:138-151) */
int cohcuda_wait_block(cohcuda_t *h, cohcuda_result *r);

typedef struct {
    const uint8_t *iq;            /* [n_blocks][C+1][B] raw bytes                             */
    int            iq_on_device;  /* 1: iq is a device pointer (16-byte aligned)              */
    uint32_t       n_blocks;
    const uint8_t *lag_request;   /* host [n_blocks][C+1] flags, or NULL -> lag_mode          */
    int            lag_mode;      /* COHCUDA_LAG_NONE / COHCUDA_LAG_ALL                       */
    const uint8_t *refnoise;      /* host [n_blocks] crefnoise::isenabled() per block or NULL */
    int            refnoise_enabled; /* used when refnoise == NULL                            */
    int            input_is_signed;
    const uint32_t *readcnt;      /* host [n_blocks][C+1] or NULL -> readcnt0 + block         */
    uint32_t       readcnt0;
    uint32_t       globalseqn0;   /* hdr0.globalseqn of block 0 (cpacketizer.cc:113)          */
    int8_t        *packets;       /* out: host (contiguous, packet_len pitch) or device
                                     (dev_pitch/dev_lead layout); NULL = keep internal        */
    int            packets_on_device;
    float         *lag;           /* out, host [n_blocks][C+1] or NULL                        */
    float         *mag;           /* out, host [n_blocks][C+1] or NULL                        */
    float         *phasecorr;     /* out, host [n_blocks][C+1][2] or NULL                     */
    float         *cf32;          /* out, DEVICE [n_blocks][C+1][L][2] corrected float matrix
                                     (column 0 = converted ref) or NULL                       */
    int            gather;        /* 1: this rank's share of a sharded batch -- its packet rows /
                                     columns go to the root rank's gathered packets (see
                                     cohcuda_gather_init below); `packets` is ignored           */
    uint32_t       gather_skip_blocks; /* block shards: leading blocks of this call that are the
                                     shard's warm-up halo -- computed, not gathered             */
    int            sticky_requests; /* 1: a lag_request flag stays pending until the channel gets a
                                     correlation slot, like csdrdevice::lagrequested
                                     (include/csdrdevice.h:129-151, src/ccoherent.cc:124,266);
                                     0: flags are the per-block is_lagrequested() state as given */
} cohcuda_batch_args;

/* Offline / bench entry: the same per-block schedule over n_blocks consecutive blocks, state
 * carried from block to block exactly as a sequence of submit/wait calls would.  Enqueues on
 * the handle's stream and returns; cohcuda_sync() waits.  With host buffers the copies run on the
 * library's own copy streams, chunk by chunk around the kernels, and consecutive batches overlap
 * (the input copies of one with the packet read-back of the previous one): host outputs are complete
 * after cohcuda_sync(), or -- for work ordered on the handle's stream -- after cohcuda_flush(). */
int cohcuda_batch(cohcuda_t *h, const cohcuda_batch_args *a);
int cohcuda_sync(cohcuda_t *h);
/* stream-ordered join (no host wait): whatever is enqueued on the handle's stream after this call
 * runs after every packet read-back issued so far */
int cohcuda_flush(cohcuda_t *h);
/* kernels launched by this handle so far (bench.py reports it as gpu_launches) */
uint64_t cohcuda_launch_count(cohcuda_t *h);
/* per-kernel timing for bench.py's roofline line: while enabled every launch is bracketed by
 * CUDA events on the launching stream and synchronised (a measurement mode, not the fast path).
 * ms3 / launches3 index: 0 = k_ref_spectrum, 1 = k_correlate, 2 = k_apply */
int cohcuda_profile(cohcuda_t *h, int enable);
int cohcuda_profile_read(cohcuda_t *h, double *ms3, uint32_t *launches3);
/* device pointer of the internal packet buffer of the last batch (dev_pitch layout) */
const int8_t *cohcuda_device_packets(cohcuda_t *h);

/* Downstream of the path (no parity target in the reference program itself): the
 * sums behind the clients' sample covariance Rxx = X^H X / L (beamformclient/heatmap2d2.cpp:199,
 * matlabclient/functions/pmusic.m:12-14), exact in integers over the int8 packet columns:
 *   S_host[b][i][j] = sum_n conj(x_i[n]) x_j[n]  as int64 {re, im}, filled for i <= j (Hermitian),
 *   m_host[b][i]    = sum_n x_i[n]               as int32 {re, im}  (the clients remove the mean).
 * packets_dev: device packets in the dev_pitch layout, or NULL for the last batch's internal
 * buffer.  Synchronous. */
int cohcuda_covariance(cohcuda_t *h, const int8_t *packets_dev, uint32_t n_blocks,
                       int64_t *S_host, int32_t *m_host);

/* Fractional part of the delay of every (block, channel) of the LAST batch, by three-point
 * parabolic interpolation of |c|^2 around the integer lag that batch reported -- the estimate
 * ccoherent::computelag prepares and leaves disabled (src/ccoherent.cc:206-218), with the curvature
 * sign corrected.  The three correlation values are recomputed exactly in integers in the time
 * domain.  frac_host: [n_blocks][C+1] (entry 0 = 0), total delay = lag + frac;  y3_host (optional):
 * [n_blocks][C+1][3] = |c[lag-1]|^2, |c[lag]|^2, |c[lag+1]|^2.  iq_dev: the batch's device input, or
 * NULL when the batch was given host buffers (the library's staged copy is used).  Synchronous. */
int cohcuda_fraclag(cohcuda_t *h, const uint8_t *iq_dev, uint32_t n_blocks, int input_is_signed,
                    float *frac_host, double *y3_host);

/* ---- lag measurement without any allocation ------------------------------------------------------
 * (No counterpart in the reference: it exists for the deployment constraint described here.)
 *
 * Who needs it.  Everything above works on a handle, and cohcuda_create() / cohcuda_batch() get their
 * working memory from cudaMalloc.  Some programs that want the lag of a few blocks cannot do that: a
 * receiver's DSP thread with a block period to keep (cudaMalloc and cudaFree serialise against the whole
 * device and can stall it for longer than one 3.4 ms block); a plug-in living inside a host program
 * whose own memory pool already owns the GPU, where a late cudaMalloc fails at some arbitrary block; a
 * process whose policy is a fixed memory map with the CUDA allocator switched off.  This entry point
 * needs no handle and reaches no cudaMalloc, cudaFree or allocating library call: its working memory
 * (twiddle tables, reference spectra, peak records, the persistent kernel's ticket) is one fixed
 * `static __device__` array inside the library, about 4.2 MiB, which every GPU that loads
 * libcoherent_cuda.so holds from the process's first CUDA call until exit -- used or not.
 *
 * What it computes.  For iq_dev = [n_blocks][n_signal + 1][blocksize] raw bytes ON THE CURRENT DEVICE
 * (channel 0 = reference, 16-byte aligned) it writes lag_dev / mag_dev = [n_blocks][n_signal + 1]
 * floats, also device memory supplied by the caller: exactly the `lag` and `mag` cohcuda_batch()
 * reports with lag_mode = COHCUDA_LAG_ALL (same kernels, bit-identical; entry 0 of each block is 0).
 * The work is enqueued on `stream` (a cudaStream_t, NULL = the default stream) and the call returns;
 * the outputs are complete once that stream has been synchronised.
 *
 * Limits (the fixed array is the reason; beyond them the call fails with COHCUDA_ERR_UNSUPPORTED and
 * a message naming the limit -- use cohcuda_batch for larger jobs, or split the call):
 *   blocksize                  power of two, 64 .. COHCUDA_STATIC_MAX_BLOCKSIZE
 *   n_blocks * 8 * blocksize   <= COHCUDA_STATIC_MAX_SPECTRUM_BYTES   (32 blocks of 16384 bytes)
 *   n_blocks * (n_signal + 1)  <= COHCUDA_STATIC_MAX_ITEMS
 *
 * First call on each device.  It resolves the array's device address (cudaGetSymbolAddress), uploads
 * the twiddle tables with a synchronous copy and waits for the device.  Make that first call -- any
 * valid call will do -- during start-up, with the intended device current, and before any stream
 * capture is begun; later calls on that device only launch kernels and one 4-byte cudaMemsetAsync.
 *
 * Sharing.  A process has ONE such array per device and every call uses it, whatever its stream or
 * thread.  It keeps whatever the previous call left; the only word that must start at zero (the
 * ticket) is cleared by each call itself, on its stream, and every other word is written before it is
 * read.  Two calls whose GPU work overlaps in time (issued on different streams, from one thread or
 * several) race on the array and return garbage: put them on one stream, or order them with events.
 * Handles created by cohcuda_create() have their own buffers and are not affected.
 *
 * errbuf (optional, errbuf_len bytes) receives the message of a failed call; returns a cohcuda_status. */
#define COHCUDA_STATIC_MAX_BLOCKSIZE      16384u
#define COHCUDA_STATIC_MAX_SPECTRUM_BYTES (4u << 20)
#define COHCUDA_STATIC_MAX_ITEMS          4096u
int cohcuda_lags_static(const uint8_t *iq_dev, uint32_t n_signal, uint32_t blocksize, uint32_t n_blocks,
                        int input_is_signed, float *lag_dev, float *mag_dev, void *stream,
                        char *errbuf, size_t errbuf_len);

/* ---- multi-GPU: the gather of the corrected matrix to one rank -------------------
 * One process and one handle per GPU.  The path shards with no exchange inside the arithmetic;
 * the only collective is the gather of the packed int8 packets to the rank that feeds the
 * reference's unchanged publisher (cpacketize::send, src/cpacketizer.cc:109-129).  The gathered
 * packet is the one packet the reference would have written for ALL channels / blocks:
 * [hdr0 16][readcnt 4N][N columns x B] (include/cpacketizer.h:32-37, src/cpacketizer.cc:91-96,
 * 137-172), assembled ON THE DEVICE of the root: every rank's rows land directly at their place.
 *
 *   COHCUDA_SHARD_CHANNELS  rank r owns rank_count[r] signal channels for all blocks (its engine is
 *                           created with n_signal = rank_count[r]); the reference channel is
 *                           replicated, the root writes column 0, header and readcnt words.
 *                           readcnt, if given to the root's batch, is [n_blocks][N].
 *   COHCUDA_SHARD_BLOCKS    rank r owns rank_count[r] consecutive blocks for all channels.
 *
 * Transports (all move chunk k while chunk k+1 is being computed):
 *   COHCUDA_GATHER_NCCL       ncclSend / ncclRecv on a side stream, buffers registered with the
 *                             communicator; the root places received columns with a strided
 *                             device copy.  libnccl.so.2 is loaded at run time.
 *   COHCUDA_GATHER_PEER_COPY  the root's gathered buffer is mapped into every rank (CUDA IPC);
 *                             each rank's copy engine writes its columns in place over NVLink.
 *   COHCUDA_GATHER_PEER_STORE same mapping; k_apply itself stores its packed columns through the
 *                             peer mapping (compute and transfer in one kernel, run on
 *                             store_ctas SMs beside the next chunk's correlation).
 * Call order: gather_init on every rank; gather_export on the root; hand the blob to every rank
 * (torch.distributed / MPI / a socket -- host plumbing, not part of this library); gather_attach
 * on every rank (collective for NCCL).  Then cohcuda_batch(..., gather = 1) on every rank,
 * cohcuda_sync(), and a barrier between the ranks before the root reads the packets (for NCCL
 * the root's own cohcuda_sync() suffices).  Two gathered buffers alternate from batch to batch.
 */
#define COHCUDA_SHARD_CHANNELS 0
#define COHCUDA_SHARD_BLOCKS   1
#define COHCUDA_GATHER_NCCL       0
#define COHCUDA_GATHER_PEER_COPY  1
#define COHCUDA_GATHER_PEER_STORE 2
#define COHCUDA_GATHER_BLOB_MAX   256

typedef struct {
    int rank, n_ranks, root;
    int shard;                    /* COHCUDA_SHARD_*                                            */
    int transport;                /* COHCUDA_GATHER_*                                           */
    const uint32_t *rank_count;   /* n_ranks entries: signal channels / blocks of each rank     */
    uint32_t max_blocks;          /* blocks per gathered batch the buffers must hold            */
    uint32_t chunk_blocks;        /* blocks per transfer, 0 = default                           */
    uint32_t store_ctas;          /* PEER_STORE: CTAs of the storing k_apply, 0 = default       */
    int32_t  reserve_sms;         /* NCCL: SMs the persistent kernels leave free for NCCL's send / receive
                                     kernels; 0 = default (16), negative = none, at most half the GPU    */
} cohcuda_gather_config;

typedef struct {
    size_t   packet_len;          /* bytes of one gathered packet                               */
    size_t   dev_pitch, dev_lead; /* device layout of the gathered packets (as packet_layout)   */
    uint32_t total_channels;      /* N: columns incl. the reference                             */
    uint32_t total_blocks;        /* rows the gathered buffer holds                             */
    uint32_t first_column;        /* CHANNELS: gathered column of this rank's signal channel 1  */
    uint32_t first_block;         /* BLOCKS: gathered row of this rank's first block            */
    size_t   run_offset;          /* byte offset, from the start of a gathered packet, of the
                                     contiguous run this rank contributes to each of its rows   */
    size_t   run_bytes;           /* length of that run                                         */
} cohcuda_gather_geometry;

/* pure arithmetic (no CUDA call): where rank `rank` of cfg lands in the gathered packets of an
 * engine with the given blocksize / noheader.  The library uses the same function internally. */
int cohcuda_gather_plan(const cohcuda_gather_config *cfg, int rank, uint32_t blocksize, int noheader,
                        uint32_t n_signal_blocks_mode, cohcuda_gather_geometry *out);
int cohcuda_gather_init(cohcuda_t *h, const cohcuda_gather_config *cfg);
int cohcuda_gather_export(cohcuda_t *h, void *blob, size_t *blob_len);       /* root only       */
int cohcuda_gather_attach(cohcuda_t *h, const void *blob, size_t blob_len);
int cohcuda_gather_geometry_of(cohcuda_t *h, cohcuda_gather_geometry *out);
/* root: device pointer of the gathered packets written by the last gathered batch             */
const int8_t *cohcuda_gather_packets(cohcuda_t *h);
/* root: copies gathered packets [block0, block0 + n_blocks) to host, packet_len pitch; waits  */
int cohcuda_gather_fetch(cohcuda_t *h, int8_t *host_packets, uint32_t block0, uint32_t n_blocks);
/* payload bytes this rank has sent to (or, on the root, received from) the other ranks so far */
uint64_t cohcuda_gather_bytes(cohcuda_t *h);
void cohcuda_gather_destroy(cohcuda_t *h);

/* checkpoint / restore of the per-channel state; n = C+1 entries */
int cohcuda_get_state(cohcuda_t *h, cohcuda_chan_state *st, uint32_t n);
int cohcuda_set_state(cohcuda_t *h, const cohcuda_chan_state *st, uint32_t n);

/* ---- cdsp primitives on host buffers (include/cdsp.h:36-71), each one kernel ---------- */
/* every call copies in, runs the kernel, copies out and synchronises: a drop-in for code
 * that calls cdsp:: directly, not the fast path.  h supplies device + stream.              */
int cohcuda_cdsp_convtosigned(cohcuda_t *h, const uint8_t *in, uint8_t *out, int n);            /* cdsp.cc:21-34  */
int cohcuda_cdsp_convtofloat(cohcuda_t *h, float *out, const int8_t *s8, int n);                /* cdsp.cc:36-44  */
int cohcuda_cdsp_convto8bit(cohcuda_t *h, int8_t *out, const float *in, int n_complex);         /* cdsp.cc:51-54  */
int cohcuda_cdsp_scalarmul(cohcuda_t *h, float *out, const float *in, float s_re, float s_im, int n); /* cdsp.cc:46-49 */
int cohcuda_cdsp_conj_dotproduct(cohcuda_t *h, float *res2, const float *a, const float *b, int n);   /* cdsp.cc:61-66 */
int cohcuda_cdsp_magsquared(cohcuda_t *h, float *out, const float *in, int n);                  /* cdsp.cc:100-103 */
int cohcuda_cdsp_conjugatemul(cohcuda_t *h, float *out, const float *a, const float *b, int n); /* cdsp.cc:105-108 */
int cohcuda_cdsp_indexofmax(cohcuda_t *h, uint32_t *idx, const float *in, int n);               /* cdsp.cc:135-139 */
int cohcuda_cdsp_rms(cohcuda_t *h, float *res, const float *in, int n);                         /* cdsp.cc:68-73  */
/* batched c2c DFT, n a power of two (32..16384), sign -1 forward / +1 backward, unnormalised:
 * fftwf_plan_many_dft + fftwf_execute_dft as used at src/ccoherent.cc:87-93,174,182 */
int cohcuda_cdsp_fft(cohcuda_t *h, float *out, const float *in, int n, int howmany, int sign);  /* cdsp.cc:110-120 */

#ifdef __cplusplus
}
#endif
#endif
