/*
 * oracle/shim/volk/volk.h -- declaration-only stand-in for <volk/volk.h> (GNU Radio VOLK, not
 * vendored by the reference and absent from this image).  TEST INFRASTRUCTURE ONLY: lets the
 * reference's own translation units compile, unmodified, where they lie under /root/reference
 * (oracle/Makefile target `ref`).  Only the kernels the reference calls are declared; their
 * generic-C semantics are restated in shim_impl.cc.  The by-value scalar of
 * volk_32fc_s32fc_multiply_32fc is the pre-3.1 VOLK API the reference is written against
 * (src/cdsp.cc:47 in the reference tree).
 */
#ifndef ORACLE_SHIM_VOLK_H
#define ORACLE_SHIM_VOLK_H
#include <complex>
#include <cstddef>
#include <cstdint>
#include <cmath>

typedef std::complex<float> lv_32fc_t;

extern "C++" {
size_t volk_get_alignment(void);
void  *volk_malloc(size_t size, size_t alignment);
void   volk_free(void *p);

void volk_8i_s32f_convert_32f(float *out, const int8_t *in, const float scalar, unsigned int n);
void volk_32f_s32f_convert_8i(int8_t *out, const float *in, const float scalar, unsigned int n);
void volk_32fc_s32fc_multiply_32fc(lv_32fc_t *out, const lv_32fc_t *in, const lv_32fc_t scalar, unsigned int n);
void volk_32fc_x2_conjugate_dot_prod_32fc(lv_32fc_t *result, const lv_32fc_t *in, const lv_32fc_t *taps, unsigned int n);
void volk_32f_x2_dot_prod_32f(float *result, const float *in, const float *taps, unsigned int n);
void volk_32fc_magnitude_squared_32f(float *out, const lv_32fc_t *in, unsigned int n);
void volk_32fc_x2_multiply_conjugate_32fc(lv_32fc_t *out, const lv_32fc_t *a, const lv_32fc_t *b, unsigned int n);
void volk_32f_index_max_32u(uint32_t *target, const float *src, uint32_t n);
}
#endif
