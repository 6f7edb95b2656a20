/* Minimal stand-in for <SoapySDR/Types.h>: type names only, no arithmetic.
 * TEST INFRASTRUCTURE (oracle/_ref build) - lets the reference's unmodified
 * CaribouliteStream.cpp compile without the un-vendored SoapySDR dependency. */
#pragma once
#include <stddef.h>
