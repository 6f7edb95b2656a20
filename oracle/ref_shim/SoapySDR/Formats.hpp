/* Minimal stand-in for <SoapySDR/Formats.hpp> (oracle/_ref build only). */
#pragma once
#define SOAPY_SDR_CF64 "CF64"
#define SOAPY_SDR_CF32 "CF32"
#define SOAPY_SDR_CS32 "CS32"
#define SOAPY_SDR_CS16 "CS16"
#define SOAPY_SDR_CS12 "CS12"
#define SOAPY_SDR_CS8  "CS8"
