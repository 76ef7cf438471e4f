/*
 * oracle/shim/fftw3.h -- declaration-only stand-in for <fftw3.h> (fftw3f is not vendored by the
 * reference and absent from this image).  TEST INFRASTRUCTURE ONLY.  Declares exactly the calls
 * made at src/ccoherent.cc:66-67,87-93,107-110 and src/cdsp.cc:115,129 of the reference tree;
 * shim_impl.cc implements them over the oracle's float32 c2c transform (same conventions:
 * FFTW_FORWARD = -1, FFTW_BACKWARD = +1, both unnormalised).
 */
#ifndef ORACLE_SHIM_FFTW3_H
#define ORACLE_SHIM_FFTW3_H
#include <stddef.h>
typedef float fftwf_complex[2];
struct orc_shim_plan;
typedef struct orc_shim_plan *fftwf_plan;
#define FFTW_FORWARD  (-1)
#define FFTW_BACKWARD (+1)
#define FFTW_MEASURE  (0U)
#define FFTW_ESTIMATE (1U << 6)
#ifdef __cplusplus
extern "C" {
#endif
fftwf_complex *fftwf_alloc_complex(size_t n);
void  fftwf_free(void *p);
fftwf_plan fftwf_plan_many_dft(int rank, const int *n, int howmany, fftwf_complex *in, const int *inembed,
                               int istride, int idist, fftwf_complex *out, const int *onembed,
                               int ostride, int odist, int sign, unsigned flags);
void  fftwf_execute(const fftwf_plan p);
void  fftwf_execute_dft(const fftwf_plan p, fftwf_complex *in, fftwf_complex *out);
void  fftwf_destroy_plan(fftwf_plan p);
#ifdef __cplusplus
}
#endif
#endif
