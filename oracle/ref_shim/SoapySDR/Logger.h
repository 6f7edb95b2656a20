/* Minimal stand-in for <SoapySDR/Logger.h> (oracle/_ref build only). */
#pragma once
#include <stdarg.h>
#include <stdio.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef enum { SOAPY_SDR_FATAL = 1, SOAPY_SDR_CRITICAL, SOAPY_SDR_ERROR, SOAPY_SDR_WARNING,
               SOAPY_SDR_NOTICE, SOAPY_SDR_INFO, SOAPY_SDR_DEBUG, SOAPY_SDR_TRACE, SOAPY_SDR_SSI } SoapySDRLogLevel;
static inline void SoapySDR_logf(const SoapySDRLogLevel lvl, const char *fmt, ...) { (void)lvl; (void)fmt; }
static inline void SoapySDR_log(const SoapySDRLogLevel lvl, const char *msg) { (void)lvl; (void)msg; }
#ifdef __cplusplus
}
#endif
