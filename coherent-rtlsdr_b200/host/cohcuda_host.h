/* host/cohcuda_host.h -- helpers shared by the drop-in translation units (host/cdsp.cc,
 * host/ccoherent.cc) that replace the reference's src/cdsp.cc and src/ccoherent.cc.
 * The class declarations those files implement are the reference's own headers
 * (include/cdsp.h, include/ccoherent.h), unchanged. */
#ifndef COHCUDA_HOST_H
#define COHCUDA_HOST_H
#include <stddef.h>
#include <stdint.h>
#include "coherent_cuda.h"
#ifdef __cplusplus
extern "C" {
#endif
/* process-wide handle used by the cdsp:: statics (device from $COHCUDA_DEVICE, default 0) */
cohcuda_t *cohcuda_host_prim_handle(void);
/* replacement for fftwf_plan_many_dft(1, {n}, howmany, in, .., 1, n, out, .., 1, n, sign, 0)
 * (src/ccoherent.cc:87-93): the object an fft_scheme must point to for cdsp::fft on this path */
void *cohcuda_host_make_fft_scheme(int n, int howmany, int sign, void *in, void *out);
void cohcuda_host_destroy_fft_scheme(void *scheme);
/* 1 (default): cdsp::convtosigned's in-place form -- the XOR cbuffer::setbufferptr asks for in the USB
 * callback thread (include/common.h:121) -- is deferred to the kernels' load path; the DSP thread then
 * submits offset-binary buffers with input_is_signed = 0.  0 when COHCUDA_HOST_XOR_ON_CALL=1. */
int cohcuda_host_xor_deferred(void);
/* 1 if [p, p+n) lies inside a buffer whose in-place convtosigned was deferred, i.e. its bytes are still
 * offset-binary.  cdsp::convtofloat consults the same record and applies the pending XOR itself, so code
 * that reads such buffers through cdsp:: (console `phase`, src/console.cc:295-299) sees signed samples;
 * only code that reads the raw bytes directly needs to ask. */
int cohcuda_host_xor_pending(const void *p, size_t n);
/* drops the record of the buffer that starts at p: call it when such a buffer is freed (the record is
 * keyed by address and would otherwise outlive it; librtlsdr's ring lives as long as the device) */
void cohcuda_host_xor_forget(const void *p);
/* the engine a ccoherent instance drives (NULL if none is alive); for status displays/tests */
cohcuda_t *cohcuda_host_engine(const void *ccoherent_instance);
/* blocks the DSP thread consumed WITHOUT publishing a packet because a library call failed (the clients
 * see the gap in readcnt / globalseqn); instance NULL = all instances.  A failed call no longer aborts
 * the process: the block is dropped, 50 failures in a row rebuild the engine once, and only if that
 * fails too does the thread give up (there is no CPU path to fall back to). */
uint64_t cohcuda_host_dropped_blocks(const void *ccoherent_instance);
/* test hook: the next n blocks are treated as if cohcuda_submit_block had failed */
void cohcuda_host_inject_failures(int n);
#ifdef __cplusplus
}
#endif
#endif
