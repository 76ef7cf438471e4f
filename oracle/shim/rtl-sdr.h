/*
 * oracle/shim/rtl-sdr.h -- declaration-only stand-in for librtlsdr's <rtl-sdr.h> (tejeez fork,
 * not vendored; README.md:35-36 of the reference).  TEST INFRASTRUCTURE ONLY.  shim_impl.cc
 * supplies a hardware-free fake: no USB, rtlsdr_read_async parks until cancelled (the harness
 * feeds blocks through csdrdevice::swapbuffer itself), tuning calls succeed and
 * rtlsdr_set_sample_freq_correction_f records what ccontrol asked for.
 */
#ifndef ORACLE_SHIM_RTLSDR_H
#define ORACLE_SHIM_RTLSDR_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef struct rtlsdr_dev rtlsdr_dev_t;
typedef void (*rtlsdr_read_async_cb_t)(unsigned char *buf, uint32_t len, void *ctx);
uint32_t rtlsdr_get_device_count(void);
const char *rtlsdr_get_device_name(uint32_t index);
int rtlsdr_get_device_usb_strings(uint32_t index, char *manufact, char *product, char *serial);
int rtlsdr_get_index_by_serial(const char *serial);
int rtlsdr_open(rtlsdr_dev_t **dev, uint32_t index);
int rtlsdr_close(rtlsdr_dev_t *dev);
int rtlsdr_set_center_freq(rtlsdr_dev_t *dev, uint32_t freq);
int rtlsdr_set_sample_rate(rtlsdr_dev_t *dev, uint32_t rate);
int rtlsdr_set_agc_mode(rtlsdr_dev_t *dev, int on);
int rtlsdr_set_tuner_gain(rtlsdr_dev_t *dev, int gain);
int rtlsdr_set_tuner_gain_mode(rtlsdr_dev_t *dev, int manual);
int rtlsdr_set_dithering(rtlsdr_dev_t *dev, int dither);
int rtlsdr_set_sample_freq_correction_f(rtlsdr_dev_t *dev, float ppm);
int rtlsdr_reset_buffer(rtlsdr_dev_t *dev);
int rtlsdr_read_async(rtlsdr_dev_t *dev, rtlsdr_read_async_cb_t cb, void *ctx, uint32_t buf_num, uint32_t buf_len);
int rtlsdr_cancel_async(rtlsdr_dev_t *dev);
#ifdef __cplusplus
}
#endif
#endif
