"""Link-level proof of the drop-in: tests/dropin/_build/libdropin.so is the reference's UNMODIFIED
CaribouliteStream.cpp, cariboulite_radio.c, caribou_smi.c (minus caribou_smi_read / _write), iir1 and zf_log,
linked with --no-undefined against libcaribou_cuda through a caribou_smi_read whose body is the INTEGRATION.md
section 1 snippet (tests/dropin/Makefile).  The CPU part checks the link; the GPU part calls the reference's own
Stream::ReadSamplesGen and cariboulite_radio_read_samples symbols with the GPU underneath and compares them, bit
for bit, with the all-CPU build of the same sources (oracle/_ref)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "tests", "dropin", "_build", "libdropin.so")
_p = C.c_void_p


def _need():
    if not os.path.exists(SO):
        if os.path.isdir("/root/reference"):
            subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "dropin")])
        else:
            pytest.skip("tests/dropin/_build/libdropin.so is built where /root/reference exists and travels with the snapshot")


def test_dropin_links_the_unmodified_reference_against_libcaribou_cuda(cb):
    _need()
    syms = subprocess.check_output(["nm", "-D", "--defined-only", SO], text=True)
    for s in ("caribou_smi_read", "caribou_smi_write", "caribou_smi_init", "caribou_smi_get_native_batch_samples",
              "cariboulite_radio_read_samples", "cariboulite_radio_get_native_mtu_size_samples",
              "_ZN8SoapySDR6Stream14ReadSamplesGenEPvml", "_ZN8SoapySDR6Stream9setFormatERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEE"):
        assert (" T " + s + "\n") in syms or (" T " + s) in syms, s
    # the reference's CPU decode is still in the object (renamed away), the exported caribou_smi_read is the GPU one
    assert " T caribou_smi_read_cpu" in syms
    und = subprocess.check_output(["nm", "-D", "--undefined-only", SO], text=True)
    for s in ("caribou_cuda_smi_open", "caribou_cuda_smi_read", "caribou_cuda_smi_write", "caribou_cuda_smi_close"):
        assert s in und, s
    assert "libcaribou_cuda.so" in subprocess.check_output(["ldd", SO], text=True)
    # generated hardware stubs only cover driver entry points (control plane)
    stubs = open(os.path.join(os.path.dirname(SO), "hw_stubs.c")).read()
    assert "caribou_smi" not in stubs and "cariboulite_radio" not in stubs


class Drop:
    def __init__(self, raw, channel=0, nbl=524288):
        L = C.CDLL(SO)
        L.dropin_dev_open.restype = _p
        L.dropin_dev_open.argtypes = [_p, C.c_size_t, C.c_size_t, C.c_int, C.c_int]
        L.dropin_dev_close.argtypes = [_p]
        L.dropin_radio_read.restype = C.c_int
        L.dropin_radio_read.argtypes = [_p, _p, _p, C.c_size_t]
        L.dropin_radio_mtu.restype = C.c_size_t
        L.dropin_radio_mtu.argtypes = [_p]
        L.dropin_stream_open.argtypes = [_p]
        L.dropin_stream_set_format.restype = C.c_int
        L.dropin_stream_set_format.argtypes = [_p, C.c_char_p]
        L.dropin_stream_set_filter.argtypes = [_p, C.c_int]
        L.dropin_stream_mtu.restype = C.c_size_t
        L.dropin_stream_mtu.argtypes = [_p]
        L.dropin_stream_read.restype = C.c_int
        L.dropin_stream_read.argtypes = [_p, _p, C.c_size_t]
        self.L = L
        raw = np.ascontiguousarray(raw).view(np.uint8).reshape(-1)
        self.h = L.dropin_dev_open(raw.ctypes.data, raw.size, nbl, channel, 0)
        assert self.h, "dropin_dev_open failed (needs a B200: no CPU fallback)"
        self.fmt = "CS16"

    def close(self):
        if self.h:
            self.L.dropin_dev_close(self.h)
            self.h = None

    def stream_read(self, n):
        dt = {"CS16": np.int16, "CS8": np.int8, "CF32": np.float32, "CF64": np.float64}[self.fmt]
        buf = np.zeros((n + 1, 2), dtype=dt)
        ret = self.L.dropin_stream_read(self.h, buf.ctypes.data, n)
        return ret, buf[:max(ret, 0)]


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", ["CS16", "CS8", "CF32", "CF64"])
@pytest.mark.parametrize("filt", [0, 1, 2, 3])
def test_reference_stream_reads_through_the_gpu(cb, po, torch_cuda, fmt, filt):
    """The 16 format x filter combinations of SoapySDR::Stream::ReadSamplesGen (LC/soapy_api/CaribouliteStream.cpp:370-382),
    the reference's own object code on both sides; the only difference is what caribou_smi_read runs on.  A
    misaligned capture (3 bytes dropped), so the extrapolated samples and the stale slots of the Stream's persistent
    buffer are in play.  Bit-exact."""
    _need()
    if not po.ref_available():
        pytest.skip("oracle/_ref is missing")
    raw = np.ascontiguousarray(po.gen_words(1 << 19).view(np.uint8)[3:])
    ref = po.RefDev(raw)
    ref.stream_open()
    d = Drop(raw)
    try:
        d.L.dropin_stream_open(d.h)
        assert d.L.dropin_stream_mtu(d.h) == 131072 == d.L.dropin_radio_mtu(d.h)
        assert d.L.dropin_stream_set_format(d.h, fmt.encode()) == ref.stream_set_format(fmt) == 0
        d.fmt = fmt
        d.L.dropin_stream_set_filter(d.h, filt)
        ref.stream_set_filter(filt)
        for req in [1000, 131072, 200000, 50000, 131072]:
            r, got = d.stream_read(req)
            rr, exp = ref.stream_read(req)
            assert r == rr, (req, r, rr)
            np.testing.assert_array_equal(got, exp)
    finally:
        d.close()
        ref.close()


@pytest.mark.gpu
@pytest.mark.parametrize("channel", [0, 1])
def test_reference_radio_read_through_the_gpu(cb, po, torch_cuda, channel):
    """cariboulite_radio_read_samples (LC/cariboulite_radio.c:1258-1285, unmodified) -> GPU caribou_smi_read: return
    codes (incl. -3 and the timeout 0 at end of data), CS16 and metadata with sentinels in the untouched slots."""
    _need()
    if not po.ref_available():
        pytest.skip("oracle/_ref is missing")
    n = 1 << 19
    w = po.gen_words(n, channel=channel).copy()
    w[3 * 131072:3 * 131072 + 131072] = 0                 # chunk 3: no alignment anywhere -> -3
    raw = np.ascontiguousarray(w.view(np.uint8)[5:])
    ref = po.RefDev(raw, channel)
    d = Drop(raw, channel)
    try:
        seen = []
        for req in [131072, 131072, 131072, 131072, 100000, 131072]:
            cs16 = np.full((req + 1, 2), 0x1234, dtype=np.int16)
            meta = np.full(req + 1, 0x34, dtype=np.uint8)          # RefDev.smi_read fills meta with sentinel & 0xFF
            r = d.L.dropin_radio_read(d.h, cs16.ctypes.data, meta.ctypes.data, req)
            rr, ecs, emeta = ref.smi_read(req, sentinel=0x1234)
            assert r == rr, (req, r, rr)
            seen.append(r)
            np.testing.assert_array_equal(cs16, ecs)
            np.testing.assert_array_equal(meta, emeta)
        assert -3 in seen and 0 in seen
    finally:
        d.close()
        ref.close()
