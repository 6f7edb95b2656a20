"""Drop-in shims (include/caribou_cuda.h "drop-in shims"): the caribou_smi_read-shaped entry point and the RX half
of SoapySDR::Stream, fed from a memfd like the reference build of the oracle is, compared with the oracle call by
call (return codes included).  Also the device-side synthetic generator against the host one."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _memfd(raw):
    fd = os.memfd_create("smi")
    b = np.ascontiguousarray(raw).view(np.uint8).tobytes()
    os.write(fd, b)
    os.lseek(fd, 0, os.SEEK_SET)
    return fd


class Smi:
    def __init__(self, cb, raw, nbl=524288):
        self.cb, self.L = cb, cb.lib()
        self.fd = _memfd(raw)
        self.h = self.L.caribou_cuda_smi_open(self.fd, nbl, 0)
        assert self.h, cb.last_error()

    def read(self, n, channel=0, sentinel=0x1234, want_cs16=True, want_meta=True):
        cs16 = np.full((n + 1, 2), sentinel, dtype=np.int16)
        meta = np.full(n + 1, 0x5A, dtype=np.uint8)
        r = self.L.caribou_cuda_smi_read(self.h, channel, cs16.ctypes.data if want_cs16 else None,
                                         meta.ctypes.data if want_meta else None, n)
        return r, cs16, meta

    def close(self):
        self.L.caribou_cuda_smi_close(self.h)
        os.close(self.fd)


@pytest.mark.parametrize("channel", [0, 1])
@pytest.mark.parametrize("drop", [0, 3, 5])
def test_smi_read_shim_matches_oracle(cb, po, torch_cuda, channel, drop):
    """int caribou_smi_read(...) (caribou_smi.h:96-97): sequence of reads over one capture, misaligned included;
    untouched slots keep the caller's bytes; 0 at end of data."""
    n = 1 << 19
    raw = np.ascontiguousarray(po.gen_words(n, channel=channel).view(np.uint8)[drop:])
    s = Smi(cb, raw)
    cursor = 0
    try:
        for req in [1000, 131072, 200000, 131072 * 2, 7, 400000]:
            got_r, got_cs16, got_meta = s.read(req, channel)
            ref_cs16 = np.full((req + 1, 2), 0x1234, dtype=np.int16)
            ref_meta = np.full(req + 1, 0x5A, dtype=np.uint8)
            ref_r, _, _, cursor = po.smi_read(raw, req, channel, 524288, cursor, cs16=ref_cs16, meta=ref_meta)
            assert got_r == ref_r, (req, got_r, ref_r)
            np.testing.assert_array_equal(got_cs16, ref_cs16)
            np.testing.assert_array_equal(got_meta, ref_meta)
        assert s.L.caribou_cuda_smi_get_native_batch_samples(s.h) == 131072
    finally:
        s.close()


def test_smi_read_shim_sync_lost_and_null_buffers(cb, po, torch_cuda):
    n = 1 << 18
    w = po.gen_words(n).copy()
    w[131072:131072 * 2] = 0                    # a whole chunk without any alignable header: -3 (caribou_smi.c:665-668)
    raw = w.view(np.uint8)
    s = Smi(cb, raw)
    try:
        r, cs16, meta = s.read(131072, want_meta=False)
        ref_r, ref_cs16, _, cur = po.smi_read(raw, 131072, 0)
        assert r == ref_r == 131072
        np.testing.assert_array_equal(cs16[:131072], ref_cs16[:131072])
        assert np.all(meta == 0x5A)             # metadata == NULL: never written
        r, _, _ = s.read(131072)
        ref_r, _, _, _ = po.smi_read(raw, 131072, 0, cursor=cur)
        assert r == ref_r == -3
    finally:
        s.close()


FMT = {"CS16": (np.int16, 4), "CS8": (np.int8, 2), "CF32": (np.float32, 8), "CF64": (np.float64, 16)}


@pytest.mark.parametrize("fmt", ["CS16", "CF32", "CF64", "CS8"])
@pytest.mark.parametrize("filt", [0, 1, 2, 3])
def test_rxstream_shim_matches_stream_oracle(cb, po, torch_cuda, fmt, filt):
    """Stream::ReadSamplesGen (CaribouliteStream.cpp:370-382): format switch, MTU clamp for non-CS16, persistent
    filter state across calls, every filter the Stream ships (type 1 = fc 10 kHz is the narrowest: look-back
    window 6).  Unfiltered: bit-exact.  Filtered (int16 truncation after the filter): at most one LSB on fewer
    than 1e-4 of the samples, and 1e-5 of signal RMS per rail over the whole stream (north_star tolerance)."""
    L = cb.lib()
    raw = np.ascontiguousarray(po.gen_words(1 << 19).view(np.uint8)[3:])
    so = po.StreamOracle(raw)
    so.set_format(fmt)
    so.set_filter(filt)
    s = Smi(cb, raw)
    st = L.caribou_cuda_rxstream_create(s.h, 0)
    assert st, cb.last_error()
    all_got, all_ref = [], []
    try:
        assert L.caribou_cuda_rxstream_set_format(st, b"XX") == -1
        assert L.caribou_cuda_rxstream_set_format(st, fmt.encode()) == 0
        L.caribou_cuda_rxstream_set_filter(st, filt)
        assert L.caribou_cuda_rxstream_mtu(st) == 131072
        dt, _ = FMT[fmt]
        for req in [1000, 131072, 200000, 50000, 131072]:
            buf = np.zeros((req + 1, 2), dtype=dt)
            r = L.caribou_cuda_rxstream_read(st, buf.ctypes.data, req, 100000)
            ref_r, ref = so.read(req)
            assert r == ref_r, (req, r, ref_r)
            got = buf[:r]
            assert np.all(buf[r:] == 0)              # nothing beyond what was delivered
            if filt == 0:
                np.testing.assert_array_equal(got, ref)
            else:
                lsb = {"CS16": 1, "CS8": 1, "CF32": 1 / 4096, "CF64": 1 / 4096}[fmt]
                d = np.abs(got.astype(np.float64) - ref.astype(np.float64))
                assert d.max() <= lsb * 1.0001
                assert np.mean(d > 0) < 1e-4
            all_got.append(got.astype(np.float64))
            all_ref.append(ref.astype(np.float64))
    finally:
        L.caribou_cuda_rxstream_destroy(st)
        s.close()
    g, r = np.concatenate(all_got), np.concatenate(all_ref)
    if fmt != "CS8":                                  # 8-bit output: one LSB is 2^-7 of full scale, RMS says nothing
        err = np.sqrt(np.mean((g - r) ** 2, axis=0)) / np.sqrt(np.mean(r ** 2, axis=0))
        assert np.all(err <= 1e-5), err


def test_rxstream_shim_lost_sync_returns_nothing_and_keeps_the_filters(cb, po, torch_cuda):
    """A read whose chunk has no alignment returns 0 elements (CaribouliteStream.cpp:266-276) and never reaches the
    filter loop: the next good read continues from the state before it, like the reference's never-reset filters."""
    L = cb.lib()
    n = 131072
    w = po.gen_words(3 * n).copy()
    w[n:2 * n] = 0
    raw = w.view(np.uint8)
    so = po.StreamOracle(raw)
    so.set_format("CF32")
    so.set_filter(2)
    s = Smi(cb, raw)
    st = L.caribou_cuda_rxstream_create(s.h, 0)
    try:
        L.caribou_cuda_rxstream_set_format(st, b"CF32")
        L.caribou_cuda_rxstream_set_filter(st, 2)
        for k in range(3):
            buf = np.full((n, 2), 9.0, dtype=np.float32)
            r = L.caribou_cuda_rxstream_read(st, buf.ctypes.data, n, 100000)
            ref_r, ref = so.read(n)
            assert r == ref_r == (0 if k == 1 else n)
            if k == 1:
                assert np.all(buf == 9.0)
            else:
                d = np.abs(buf.astype(np.float64) - ref.astype(np.float64))
                assert d.max() <= 1.0001 / 4096 and np.mean(d > 0) < 1e-4
    finally:
        L.caribou_cuda_rxstream_destroy(st)
        s.close()


@pytest.mark.parametrize("channel", [0, 1])
@pytest.mark.parametrize("drop", [0, 5])
def test_radio_read_cf32_is_the_cpp_api_read(cb, po, torch_cuda, channel, drop):
    """CaribouLiteRadio::ReadSamples(std::complex<float>*, n, meta) (LC/CaribouLiteRadioCpp.cpp:72-130): the native read
    lands in the radio's persistent buffers, then samples = native / 4096.0f and the metadata is copied out - so a
    slot a misaligned chunk leaves unwritten carries the PREVIOUS read's value.  Oracle: the reference's read loop
    over persistent arrays, converted."""
    L = cb.lib()
    n = 1 << 19
    raw = np.ascontiguousarray(po.gen_words(n, channel=channel).view(np.uint8)[drop:])
    s = Smi(cb, raw)
    cap = 131072 * 2
    p_cs16 = np.zeros((cap + 1, 2), dtype=np.int16)      # _read_samples / _read_metadata
    p_meta = np.zeros(cap + 1, dtype=np.uint8)
    cursor = 0
    try:
        for req in [131072 * 2, 131072, 70000, 131072 * 2]:
            cf = np.full((req, 2), 5.0, dtype=np.float32)
            meta = np.full(req, 0x77, dtype=np.uint8)
            r = L.caribou_cuda_radio_read_cf32(s.h, channel, cf.ctypes.data, meta.ctypes.data, req)
            ref_r, _, _, cursor = po.smi_read(raw, req, channel, 524288, cursor, cs16=p_cs16, meta=p_meta)
            assert r == ref_r, (req, r, ref_r)
            if r > 0:
                np.testing.assert_array_equal(cf[:r], po.cs16_to_cf32(p_cs16[:r]))
                np.testing.assert_array_equal(meta[:r], p_meta[:r])
            assert np.all(cf[max(r, 0):] == 5.0) and np.all(meta[max(r, 0):] == 0x77)
    finally:
        s.close()


@pytest.mark.parametrize("hif", [0, 1])
def test_device_generator_matches_host_generator(cb, po, torch_cuda, hif):
    """the device-side synthetic word generator is cross-checked against the oracle's on a 2^22
    prefix and on a window far into the stream (64-bit sample index)."""
    torch = torch_cuda
    ti, tq = po.tone_tables()
    with cb.Context() as ctx:
        for n0, n, seed in [(0, 1 << 22, 0xC0FFEE + hif), ((1 << 33) - 4096, 1 << 16, 12345)]:
            d = torch.zeros(n, dtype=torch.int32, device="cuda:0")
            ctx.synth_words(d, n, n0=n0, hif_layout=hif, seed=seed, tone_i=ti, tone_q=tq)
            ctx.wait()
            ref = po.gen_words(n, n0=n0, channel=hif, seed=seed)
            np.testing.assert_array_equal(d.cpu().numpy().view(np.uint32), ref)


def _out_fd():
    return os.memfd_create("out")


def _read_all(fd, dtype):
    size = os.lseek(fd, 0, os.SEEK_END)
    os.lseek(fd, 0, os.SEEK_SET)
    return np.frombuffer(os.read(fd, size), dtype=dtype).copy()


def test_pump_words_to_cs16_is_the_cariboulite_util_capture(cb, po, torch_cuda):
    """caribou_cuda_pump without filter = the capture loop of cariboulite_util (cariboulite_util.c:340-361): the
    output file is the raw little-endian CS16 the reference would have written, bit for bit (ragged last block)."""
    n = 5 * 131072 + 4000
    w = po.gen_words(n)
    fin, fout = _memfd(w), _out_fd()
    with cb.Context() as ctx:
        r = ctx.pump(cb.IN_WORDS, fin, fout, 2 * 131072)
    assert r == n
    got = _read_all(fout, np.int16).reshape(-1, 2)
    ref_ret, ref_cs16, _, _ = po.smi_read(w.view(np.uint8), n, 0)
    assert ref_ret == n
    np.testing.assert_array_equal(got, ref_cs16[:n])
    os.close(fin); os.close(fout)


@pytest.mark.parametrize("in_kind", ["words", "cs16"])
def test_pump_filtered_decimated_stream(cb, po, torch_cuda, in_kind):
    """Filtered pump over several blocks (state carries from block to block), max_samples honoured."""
    from tests.test_gpu_iir import rel_rms, TOL
    n, D, limit = 6 * 131072, 8, 5 * 131072
    sos = po.butter_lowpass_sos(6, 4e6, 25e3)
    w = po.gen_words(n)
    _, cs16, _, _ = po.smi_read(w.view(np.uint8), n, 0)
    ref = po.filter_cf32_f32(sos, po.cs16_to_cf32(cs16[:limit]))[::D]
    fin = _memfd(w if in_kind == "words" else cs16[:n])
    fout = _out_fd()
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32) as ctx:
        r = ctx.pump(cb.IN_WORDS if in_kind == "words" else cb.IN_CS16, fin, fout, 2 * 131072, max_samples=limit)
    assert r == limit
    got = _read_all(fout, np.float32).reshape(-1, 2)
    assert got.shape[0] == limit // D
    assert np.all(rel_rms(got, ref) <= TOL)
    os.close(fin); os.close(fout)


def test_pump_reports_lost_sync(cb, po, torch_cuda):
    n = 4 * 131072
    w = po.gen_words(n).copy()
    w[2 * 131072:3 * 131072] = 0
    fin, fout = _memfd(w), _out_fd()
    with cb.Context() as ctx:
        r = ctx.pump(cb.IN_WORDS, fin, fout, 131072)
    assert r == cb.ERR_SYNC
    got = _read_all(fout, np.int16).reshape(-1, 2)
    _, ref_cs16, _, _ = po.smi_read(w.view(np.uint8), 2 * 131072, 0)
    np.testing.assert_array_equal(got, ref_cs16[:2 * 131072])       # the blocks before the bad one were delivered
    os.close(fin); os.close(fout)


def _oracle_mtu_reads(po, raw, mtu, channel=0):
    """What the reader thread queues: caribou_smi_read of one MTU at a time until the capture runs dry."""
    out, cursor = [], 0
    buf = np.zeros((mtu + 1, 2), dtype=np.int16)        # one buffer reused like interm_native_buffer1: slots a
    meta = np.zeros(mtu + 1, dtype=np.uint8)            # misaligned chunk leaves untouched keep the previous block
    while True:
        r, _, _, cursor = po.smi_read(raw, mtu, channel, mtu * 4, cursor, cs16=buf, meta=meta)
        if r <= 0:
            break
        out.append(buf[:r].copy())
    return np.concatenate(out)


@pytest.mark.parametrize("drop", [0, 3])
def test_async_reader_delivers_the_stream_in_order(cb, po, torch_cuda, drop):
    """ReaderThread + rx_queue->get (CaribouliteStream.cpp:16-49, 263): inactive until activated, then every MTU the
    GPU path decodes arrives once and in order; a read that cannot be satisfied in time takes nothing."""
    import time
    L = cb.lib()
    mtu = 8192
    raw = np.ascontiguousarray(po.gen_words(mtu * 7 + 100).view(np.uint8)[drop:])
    ref = _oracle_mtu_reads(po, raw, mtu)
    s = Smi(cb, raw, nbl=mtu * 4)
    q = L.caribou_cuda_async_open(s.h, 0, 0)
    assert q, cb.last_error()
    try:
        ring = L.caribou_cuda_async_ring(q)
        assert L.caribou_cuda_ring_capacity(ring) == 131072                 # 10 MTUs rounded up to a power of two
        buf = np.zeros((4096, 2), dtype=np.int16)
        assert L.caribou_cuda_async_read(q, buf.ctypes.data, 100, 30000) == 0     # not active yet
        L.caribou_cuda_async_activate(q, 1)
        got, total = [], 0
        for n in [100, 4096, 1, 3000, 4096, 4096, 4096] * 4:
            n = min(n, ref.shape[0] - total)
            if n == 0:
                break
            assert L.caribou_cuda_async_read(q, buf.ctypes.data, n, 5000000) == n
            got.append(buf[:n].copy()); total += n
        np.testing.assert_array_equal(np.concatenate(got), ref)
        t0 = time.time()
        while not L.caribou_cuda_async_idle(q) and time.time() - t0 < 5:
            time.sleep(0.01)
        assert L.caribou_cuda_async_idle(q) == 1
        assert L.caribou_cuda_async_read(q, buf.ctypes.data, 1, 20000) == 0       # the capture ran dry
    finally:
        L.caribou_cuda_async_close(q)
        s.close()


def test_async_reader_overrides_the_oldest_samples(cb, po, torch_cuda):
    """USE_ASYNC_OVERRIDE_WRITES (CaribouliteStream.cpp:11): a consumer that does not read loses the oldest samples,
    never the newest."""
    import time
    L = cb.lib()
    mtu = 8192
    raw = po.gen_words(mtu * 6).view(np.uint8)
    ref = _oracle_mtu_reads(po, raw, mtu)
    s = Smi(cb, raw, nbl=mtu * 4)
    q = L.caribou_cuda_async_open(s.h, 0, 2 * mtu)
    assert q, cb.last_error()
    try:
        L.caribou_cuda_async_activate(q, 1)
        t0 = time.time()
        while not L.caribou_cuda_async_idle(q) and time.time() - t0 < 10:
            time.sleep(0.01)
        ring = L.caribou_cuda_async_ring(q)
        assert L.caribou_cuda_ring_size(ring) == 2 * mtu
        assert L.caribou_cuda_ring_dropped(ring) == 4 * mtu
        buf = np.zeros((2 * mtu, 2), dtype=np.int16)
        assert L.caribou_cuda_async_read(q, buf.ctypes.data, 2 * mtu, 100000) == 2 * mtu
        np.testing.assert_array_equal(buf, ref[-2 * mtu:])
    finally:
        L.caribou_cuda_async_close(q)
        s.close()
