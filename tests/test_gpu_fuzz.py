"""Seeded random configurations of the fused rx path against the CPU oracle: filter order, decimation, channel,
number of captures, capture lengths around run / tile / chunk boundaries, one to three submits queued before the wait
(state and decimation phase carried), device or page-locked host memory, and now and then a misaligned or un-alignable
chunk in one capture.  Every kernel on the path sees shapes the hand-written cases do not list: the control-block
preparation, the alignment search over partial CTAs, the scan / rerun kernels with partial tiles and gated captures,
the tail kernel as remainder walker and as commit, the slow path for the gated capture.  Runs through the C ABI."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-5
CHUNK = 16384            # bytes: 4096 samples per chunk
SEED = int(os.environ.get("CARIBOU_FUZZ_SEED", "0"))      # shifts every case's seed: other draws of the same distributions


def rel_rms(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    d = np.sqrt(np.mean((a - b) ** 2, axis=0))
    s = np.sqrt(np.mean(b ** 2, axis=0))
    return d / np.maximum(s, 1e-30)


def _oracle_calls(po, sos, D, channel, calls):
    """One capture over several reads as the reference delivers it (test_gpu_batch.py: caribou_smi_read into a zeroed
    buffer, the never-reset filter pair over what the read returned; on -3 the chunks before the un-alignable one)."""
    ci, cq = po.Cascade(sos), po.Cascade(sos)
    outs = []
    for w in calls:
        n = w.size
        ret, cs16, _, _ = po.smi_read(w.view(np.uint8), n, channel, CHUNK)
        deliver = n
        if ret == -3:
            spc = CHUNK // 4
            nch = (n + spc - 1) // spc
            bad = next(c for c in range(nch) if po.find_buffer_offset(w[c * spc:(c + 1) * spc].view(np.uint8)) < 0)
            deliver = bad * spc
        deliver -= deliver % D
        outs.append((ret, deliver, po.filter_cf32_f32(sos, po.cs16_to_cf32(cs16[:deliver]), ci, cq)[::D]))
    return outs


CASES = list(range(72))


@pytest.mark.parametrize("case", CASES)
def test_random_rx_configurations(cb, po, torch_cuda, case):
    torch = torch_cuda
    rng = np.random.default_rng(1000 + case + 100000 * SEED)
    order, fc = [(2, 100e3), (3, 50e3), (4, 250e3), (6, 10e3), (6, 50e3), (5, 100e3), (8, 400e3)][rng.integers(7)]
    D = int([1, 1, 2, 8, 8, 4, 16][rng.integers(7)])
    channel = int(rng.integers(2))
    ns = int([1, 1, 2, 3, 5][rng.integers(5)])
    n_calls = int(rng.integers(1, 4))
    pinned = bool(rng.integers(2))
    # lengths: multiples of 32 * D (whole words, decimation phase 0 at every submit), from shorter than a run to a
    # few tiles, not multiples of the chunk in general
    unit = 32 * D
    n = unit * int(rng.integers(1, max(2, (3 * 32 * 256 + 5000) // unit)))
    if rng.integers(4) == 0:
        n = unit * int(rng.integers(1, 6))                     # tiny: no whole run at all
    sos = po.butter_lowpass_sos(order, 4e6, fc)
    ws = [[po.gen_words(n, n0=c * n, channel=channel, seed=case * 100 + s) for s in range(ns)] for c in range(n_calls)]
    spc = CHUNK // 4
    flaw = int(rng.integers(4))                                 # 0, 1: clean; 2: a misaligned chunk; 3: an un-alignable one
    fs, fcall = int(rng.integers(ns)), int(rng.integers(n_calls))
    nch = (n + spc - 1) // spc
    fch = int(rng.integers(nch))
    lo, hi = fch * spc, min((fch + 1) * spc, n)
    if flaw == 2 and hi - lo > 64:
        ws[fcall][fs][lo] ^= 0x80000000                          # the chunk starts with a bad header: offs = 4
    elif flaw == 3 and hi - lo > 64:
        ws[fcall][fs][lo:hi] = 0                                 # no valid header anywhere in the chunk
    else:
        flaw = 0
    stride = (n * 4 + 15) & ~15
    nout = n // D
    bufs, outs = [], []
    for c in range(n_calls):
        raw = np.zeros((ns, stride // 4), np.uint32)
        for s in range(ns):
            raw[s, :n] = ws[c][s]
        t = torch.from_numpy(raw.view(np.int32))
        o = torch.full((ns, nout, 2), 7.0, dtype=torch.float32)
        if pinned:
            bufs.append(t.pin_memory()); outs.append(o.pin_memory())
        else:
            bufs.append(t.to("cuda:0")); outs.append(o.to("cuda:0"))
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32, n_streams=ns, chunk_bytes=CHUNK, channel=channel) as ctx:
        for c in range(n_calls):
            ctx.rx_submit(bufs[c], n * 4, outs[c], stride_bytes=stride, filt_stride=nout,
                          mem=cb.MEM_HOST_PINNED if pinned else cb.MEM_DEVICE)
        ret, res = ctx.wait()
        sr = ctx.stream_results()
    got = [o.cpu().numpy() if not pinned else o.numpy() for o in outs]
    what = dict(case=case, order=order, fc=fc, D=D, channel=channel, ns=ns, n=n, calls=n_calls, pinned=pinned, flaw=flaw,
                at=(fs, fcall, fch))
    for s in range(ns):
        ref = _oracle_calls(po, sos, D, channel, [ws[c][s] for c in range(n_calls)])
        for c in range(n_calls):
            r_ret, deliver, y = ref[c]
            m = deliver // D
            if m:
                assert np.all(rel_rms(got[c][s, :m], y) <= TOL), (what, s, c)
            if r_ret != -3:
                assert m == nout, (what, s, c)
        # the per-capture verdict of the LAST submit is what stream_results reports
        last_ret = ref[-1][0]
        assert sr[s].n_samples == (cb.ERR_SYNC if last_ret == -3 else n), (what, s)
    any_lost_last = any(_oracle_calls(po, sos, D, channel, [ws[c][s] for c in range(n_calls)])[-1][0] == -3 for s in range(ns))
    assert ret == (cb.ERR_SYNC if any_lost_last else n), what


@pytest.mark.parametrize("case", list(range(40)))
def test_random_filter_configurations(cb, po, torch_cuda, case):
    """caribou_cuda_filter_submit on CS16 / CF32 device input: both parity modes, every output format of the reference's
    ordering, random decimation and length, the stream cut into one to three calls (carried state), end state against
    the oracle's cascades."""
    torch = torch_cuda
    rng = np.random.default_rng(5000 + case + 100000 * SEED)
    order, fc = [(1, 100e3), (2, 100e3), (3, 50e3), (4, 250e3), (6, 10e3), (6, 25e3), (7, 200e3), (8, 400e3)][rng.integers(8)]
    D = int([1, 1, 2, 4, 8, 32][rng.integers(6)])
    q16 = bool(rng.integers(2))
    kind = "cs16" if q16 else ["cs16", "cf32"][rng.integers(2)]
    fmt = ["CS16", "CF32", "CF64", "CS8"][rng.integers(4)] if q16 else ["CF32", "CF64"][rng.integers(2)]
    n_calls = int(rng.integers(1, 4))
    unit = 32 * D
    per = unit * int(rng.integers(1, max(2, 40000 // unit)))
    n = per * n_calls
    sos = po.butter_lowpass_sos(order, 4e6, fc)
    _, cs16, _, _ = po.smi_read(po.gen_words(n, seed=case).view(np.uint8), n)
    cs16 = cs16[:n]
    ci, cq = po.Cascade(sos), po.Cascade(sos)
    code = {"CS16": cb.FMT_CS16, "CF32": cb.FMT_CF32, "CF64": cb.FMT_CF64, "CS8": cb.FMT_CS8}[fmt]
    if q16:
        q = po.filter_cs16_q16(sos, cs16, ci, cq)[::D]
        ref = {"CS16": q, "CF32": po.cs16_to_cf32(q), "CF64": po.cs16_to_cf64(q), "CS8": po.cs16_to_cs8(q)}[fmt]
    else:
        y = po.filter_cf32_f32(sos, po.cs16_to_cf32(cs16), ci, cq)[::D]
        ref = y if fmt == "CF32" else None
    x_dev = _dev_t(torch, cs16 if kind == "cs16" else po.cs16_to_cf32(cs16))
    out = torch.zeros((n // D, 2), dtype=getattr(torch, np.dtype(cb.FMT_DTYPE[code]).name), device="cuda:0")
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_Q16 if q16 else cb.PARITY_F32, out_format=code) as ctx:
        for c in range(n_calls):
            ctx.filter_submit(cb.IN_CS16 if kind == "cs16" else cb.IN_CF32, x_dev[c * per:(c + 1) * per], per,
                              out[c * per // D:(c + 1) * per // D])
        assert ctx.wait()[0] == per
        state = ctx.get_state()
    got = out.cpu().numpy()
    what = dict(case=case, order=order, fc=fc, D=D, q16=q16, kind=kind, fmt=fmt, calls=n_calls, per=per)
    if q16:
        diff = np.abs(got.astype(np.float64) - ref.astype(np.float64))
        lsb = {"CS16": 1, "CF32": 1 / 4096, "CF64": 1 / 4096, "CS8": 1}[fmt]
        assert diff.max() <= lsb * 1.0001 and np.mean(diff > 0) < 1e-3, what      # an LSB may flip at a truncation boundary
    elif ref is not None:
        assert np.all(rel_rms(got, ref) <= TOL), what
    else:       # CF64 of the F32 ordering: the unrounded doubles; their float image is the CF32 output
        assert np.all(rel_rms(got, y) <= TOL), what
    # end state in the reference's (v1, v2) coordinates; Q16 feeds int16-scale samples, F32 samples / 4096
    st = np.concatenate([ci.state, cq.state])
    np.testing.assert_allclose(state, st, rtol=1e-7, atol=1e-9 * max(1.0, np.abs(st).max()), err_msg=str(what))


def _dev_t(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).to("cuda:0")


@pytest.mark.parametrize("case", list(range(40)))
def test_random_unpack_configurations(cb, po, torch_cuda, case):
    """Kernel (a) on single-channel captures with random damage: dropped and inserted bytes (every chunk after one is
    misaligned by a different amount), corrupted headers, garbage stretches (lost sync), sync bits, ragged byte counts,
    chunk sizes from 64 bytes to 64 KiB - CS16, meta, CF32 (untouched slots included), flag counts and index lists and
    the return code, bit for bit against the oracle (tests/test_gpu_unpack.py::_check_single)."""
    from tests.test_gpu_unpack import _check_single
    rng = np.random.default_rng(9000 + case + 100000 * SEED)
    channel = int(rng.integers(2))
    chunk = int([64, 256, 4096, 16384, 65536][rng.integers(5)])
    n = int(rng.integers(1, 60000 if chunk >= 4096 else 6000))
    w = po.gen_words(n, channel=channel, seed=case)
    w[rng.integers(0, n, size=max(1, n // 500))] |= 1                          # sync bits
    for _ in range(int(rng.integers(0, 4))):                                    # corrupted headers
        w[int(rng.integers(n))] ^= [0x80000000, 0x40000000, 0x00008000, 0x00004000, 0x00010000][rng.integers(5)]
    if rng.integers(5) == 0 and n > 3000:                                       # a garbage stretch
        a = int(rng.integers(n - 2000))
        w[a:a + int(rng.integers(500, 2000))] = rng.integers(0, 2**32, dtype=np.uint64, size=1).astype(np.uint32)[0]
    raw = w.view(np.uint8).copy()
    for _ in range(int(rng.integers(0, 3))):                                    # dropped / inserted bytes
        p = int(rng.integers(raw.size))
        k = int(rng.integers(1, 8))
        raw = np.concatenate([raw[:p], raw[p + k:]]) if rng.integers(2) else np.concatenate([raw[:p], np.zeros(k, np.uint8), raw[p:]])
    raw = raw[: raw.size - int(rng.integers(0, 4))]                            # ragged end
    _check_single(cb, po, torch_cuda, np.ascontiguousarray(raw), channel, chunk)


@pytest.mark.parametrize("case", list(range(24)))
def test_random_smi_read_sequences(cb, po, torch_cuda, case):
    """The caribou_smi_read-shaped shim (caribou_smi.h:96-97) over a damaged capture in a memfd: a random sequence of
    read sizes (one sample to several native batches), random native batch length, both channels; return codes, CS16,
    meta and the caller's untouched bytes against the oracle's read loop, call by call, until the data runs out."""
    from tests.test_gpu_compat import Smi
    rng = np.random.default_rng(13000 + case + 100000 * SEED)
    channel = int(rng.integers(2))
    nbl = int([4096, 16384, 65536, 524288][rng.integers(4)])
    n = int(rng.integers(2000, 300000))
    w = po.gen_words(n, channel=channel, seed=case)
    w[rng.integers(0, n, size=max(1, n // 700))] |= 1
    raw = w.view(np.uint8).copy()
    for _ in range(int(rng.integers(0, 3))):
        p = int(rng.integers(raw.size))
        k = int(rng.integers(1, 6))
        raw = np.concatenate([raw[:p], raw[p + k:]])
    if rng.integers(6) == 0:
        a = int(rng.integers(max(1, raw.size - 70000)))
        raw[a:a + int(rng.integers(20000, 70000))] = 0                    # a stretch with no header: -3 somewhere
    raw = np.ascontiguousarray(raw)
    s = Smi(cb, raw, nbl=nbl)
    cursor = 0
    try:
        for step in range(12):
            req = int([1, 7, 100, nbl // 4, nbl // 4 + 1, 3 * (nbl // 4), 5000, 131072][rng.integers(8)])
            got_r, got_cs16, got_meta = s.read(req, channel)
            ref_cs16 = np.full((req + 1, 2), 0x1234, dtype=np.int16)
            ref_meta = np.full(req + 1, 0x5A, dtype=np.uint8)
            ref_r, _, _, cursor = po.smi_read(raw, req, channel, nbl, cursor, cs16=ref_cs16, meta=ref_meta)
            what = dict(case=case, step=step, req=req, nbl=nbl, channel=channel, n=n)
            assert got_r == ref_r, (what, got_r, ref_r)
            np.testing.assert_array_equal(got_cs16, ref_cs16, err_msg=str(what))
            np.testing.assert_array_equal(got_meta, ref_meta, err_msg=str(what))
            if ref_r <= 0 and ref_r != -3:
                break
    finally:
        s.close()


@pytest.mark.parametrize("case", list(range(24)))
def test_random_stream_read_sequences(cb, po, torch_cuda, case):
    """Stream::ReadSamplesGen through the shim (CaribouliteStream.cpp:370-382) over a damaged capture: random format,
    filter, native batch length and read sizes (above and below the MTU), dropped bytes and now and then a stretch
    with no header at all (a read that hits it delivers nothing and leaves the filters alone).  Call by call against
    the oracle's Stream: same element counts, unfiltered formats bit for bit, filtered ones within an LSB on a few
    samples, and beyond what was delivered exactly what the reference leaves in the caller's buffer."""
    from tests.test_gpu_compat import Smi, FMT
    L = cb.lib()
    rng = np.random.default_rng(17000 + case + 100000 * SEED)
    fmt = ["CS16", "CF32", "CF64", "CS8"][rng.integers(4)]
    filt = int(rng.integers(4))
    nbl = int([65536, 524288][rng.integers(2)])
    n = int(rng.integers(150000, 600000))
    w = po.gen_words(n, seed=case)
    raw = w.view(np.uint8).copy()
    for _ in range(int(rng.integers(0, 2))):
        p = int(rng.integers(raw.size))
        raw = np.concatenate([raw[:p], raw[p + int(rng.integers(1, 6)):]])
    if rng.integers(4) == 0:
        a = int(rng.integers(max(1, raw.size - 300000)))
        raw[a:a + int(rng.integers(nbl, 3 * nbl))] = 0
    raw = np.ascontiguousarray(raw)
    so = po.StreamOracle(raw, native_batch_len=nbl)
    so.set_format(fmt)
    so.set_filter(filt)
    s = Smi(cb, raw, nbl=nbl)
    st = L.caribou_cuda_rxstream_create(s.h, 0)
    assert st, cb.last_error()
    try:
        assert L.caribou_cuda_rxstream_set_format(st, fmt.encode()) == 0
        L.caribou_cuda_rxstream_set_filter(st, filt)
        dt, _ = FMT[fmt]
        mtu = nbl // 4
        for step in range(8):
            req = int([1, 1000, mtu - 1, mtu, mtu + 1, 2 * mtu + 17, 50000][rng.integers(7)])
            buf = np.zeros((req + 1, 2), dtype=dt)
            r = L.caribou_cuda_rxstream_read(st, buf.ctypes.data, req, 100000)
            if fmt == "CS16":
                # the native format is read straight into the caller's buffer (CaribouliteStream.cpp:282-298): chunks
                # decoded before a lost sync are in it although the call delivers nothing
                beyond = np.zeros((req + 1, 2), dtype=np.int16)
                ref_r = so._read_cs16(beyond, req)
                ref = beyond[:ref_r]
            else:
                ref_r, ref = so.read(req)
                beyond = np.zeros((req + 1, 2), dtype=dt)
            what = dict(case=case, step=step, fmt=fmt, filt=filt, nbl=nbl, req=req)
            assert r == ref_r, (what, r, ref_r)
            got = buf[:r]
            np.testing.assert_array_equal(buf[r:], beyond[r:], err_msg=str(what))
            if filt == 0:
                np.testing.assert_array_equal(got, ref, err_msg=str(what))
            elif r:
                lsb = {"CS16": 1, "CS8": 1, "CF32": 1 / 4096, "CF64": 1 / 4096}[fmt]
                dd = np.abs(got.astype(np.float64) - ref.astype(np.float64))
                assert dd.max() <= lsb * 1.0001 and np.mean(dd > 0) < 1e-3, what
    finally:
        L.caribou_cuda_rxstream_destroy(st)
        s.close()


@pytest.mark.parametrize("case", list(range(16)))
def test_random_radio_read_sequences(cb, po, torch_cuda, case):
    """CaribouLiteRadio::ReadSamples(std::complex<float>*, n, meta) through the shim (LC/CaribouLiteRadioCpp.cpp:72-130)
    over a damaged capture, random read sizes up to the radio's persistent buffer: CF32, metadata, return codes, and the
    slots a misaligned chunk leaves unwritten (they carry the previous read's value)."""
    from tests.test_gpu_compat import Smi
    L = cb.lib()
    rng = np.random.default_rng(21000 + case + 100000 * SEED)
    channel = int(rng.integers(2))
    nbl = int([16384, 65536, 524288][rng.integers(3)])
    n = int(rng.integers(50000, 500000))
    raw = po.gen_words(n, channel=channel, seed=case).view(np.uint8).copy()
    for _ in range(int(rng.integers(0, 3))):
        p = int(rng.integers(raw.size))
        raw = np.concatenate([raw[:p], raw[p + int(rng.integers(1, 6)):]])
    if rng.integers(4) == 0:
        a = int(rng.integers(max(1, raw.size - 3 * nbl)))
        raw[a:a + int(rng.integers(nbl, 2 * nbl))] = 0
    raw = np.ascontiguousarray(raw)
    s = Smi(cb, raw, nbl=nbl)
    cap = 131072 * 2
    p_cs16 = np.zeros((cap + 1, 2), dtype=np.int16)      # the radio's _read_samples / _read_metadata
    p_meta = np.zeros(cap + 1, dtype=np.uint8)
    cursor = 0
    try:
        first = True
        for step in range(8):
            req = cap if first else int([1, 777, nbl // 4, nbl // 4 + 3, 70000, cap][rng.integers(6)])
            first = False
            cf = np.full((req, 2), 5.0, dtype=np.float32)
            meta = np.full(req, 0x77, dtype=np.uint8)
            r = L.caribou_cuda_radio_read_cf32(s.h, channel, cf.ctypes.data, meta.ctypes.data, req)
            ref_r, _, _, cursor = po.smi_read(raw, req, channel, nbl, cursor, cs16=p_cs16, meta=p_meta)
            what = dict(case=case, step=step, req=req, nbl=nbl, channel=channel)
            assert r == ref_r, (what, r, ref_r)
            if r > 0:
                np.testing.assert_array_equal(cf[:r], po.cs16_to_cf32(p_cs16[:r]), err_msg=str(what))
                np.testing.assert_array_equal(meta[:r], p_meta[:r], err_msg=str(what))
            assert np.all(cf[max(r, 0):] == 5.0) and np.all(meta[max(r, 0):] == 0x77), what
            if ref_r == 0:
                break
    finally:
        s.close()


@pytest.mark.parametrize("case", list(range(24)))
def test_random_dual_channel_unpack(cb, po, torch_cuda, case):
    """Dual-channel captures with random interleave block, chunk size, length, corrupted headers and sync
    bits: each sub-stream against the oracle's demux + read, the flag counts and lists against the oracle's scan."""
    from tests.test_gpu_unpack import _unpack_dev, _oracle
    rng = np.random.default_rng(25000 + case + 100000 * SEED)
    B = int([1, 8, 64, 4096][rng.integers(4)])
    chunk = int([4096, 32768, 65536][rng.integers(3)])
    n = int(rng.integers(2 * B, 120000))
    a = po.gen_words(n, channel=0, seed=case)
    b = po.gen_words(n, channel=1, seed=case + 500)
    w = np.empty(n, np.uint32)
    blk = np.arange(n) // B % 2
    w[blk == 0] = a[: int(np.sum(blk == 0))]
    w[blk == 1] = b[: int(np.sum(blk == 1))]
    w[rng.integers(0, n, size=max(1, n // 900))] |= 1
    for _ in range(int(rng.integers(0, 4))):
        w[int(rng.integers(n))] ^= [0x80000000, 0x40000000, 0x00008000, 0x00004000][rng.integers(4)]
    g = _unpack_dev(cb, torch_cuda, w.view(np.uint8), 0, chunk, B=B)
    what = dict(case=case, B=B, chunk=chunk, n=n)
    for sub, suf in ((0, ""), (1, "_b")):
        sw = po.demux_gather(w, B, sub)
        ret, cs16, meta = _oracle(po, sw.view(np.uint8), sub, chunk)
        assert (g["ret"] if sub == 0 else g["res"].n_samples_b) == ret, (what, sub)
        k = cs16.shape[0]
        np.testing.assert_array_equal(g["cs16" + suf][:k], cs16, err_msg=str((what, sub)))      # untouched slots included
        np.testing.assert_array_equal(g["meta" + suf][:k], meta, err_msg=str((what, sub)))
        # CF32 = cs16 / 4096 wherever the reference wrote CS16; untouched slots keep the caller's bytes
        written = ~((cs16[:, 0] == np.int16(0x5A5A)) & (cs16[:, 1] == np.int16(0x5A5A)))
        exp = np.where(written[:, None], po.cs16_to_cf32(cs16), np.float32(-7.0))
        np.testing.assert_array_equal(g["cf32" + suf][:k], exp, err_msg=str((what, sub)))
    ni, inv, ns_, syn = po.flag_scan(w, cap=4096)
    assert g["res"].n_invalid == ni and g["res"].n_sync == ns_, what
    np.testing.assert_array_equal(g["inv"][: len(inv)], inv)
    np.testing.assert_array_equal(g["syn"][: len(syn)], syn)


@pytest.mark.parametrize("case", list(range(16)))
def test_random_sharded_streams(cb, po, torch_cuda, case):
    """One stream time-chunked over `world` contexts of one GPU (tests/test_gpu_shard.py: pass 1 from zero state, the
    end vectors exchanged, composition and fix-up on the device): random world, filter, decimation, shard length (from
    shorter than the fix-up window to several tiles) and number of calls, against the sequential oracle over the whole
    stream and with a look at every shard boundary; every rank ends with the same state."""
    torch = torch_cuda
    rng = np.random.default_rng(29000 + case + 100000 * SEED)
    order, fc = [(2, 100e3), (4, 250e3), (6, 10e3), (6, 50e3), (8, 400e3), (3, 50e3)][rng.integers(6)]
    D = int([1, 2, 8, 8, 4][rng.integers(5)])
    world = int(rng.integers(2, 6))
    calls = int(rng.integers(1, 4))
    unit = 32 * D
    n = unit * int(rng.integers(1, 60000 // unit))
    total = n * world * calls
    sos = po.butter_lowpass_sos(order, 4e6, fc)
    w = po.gen_words(total, seed=case)
    ref = po.rx_path_f32(w, sos, D)
    d_w = torch.from_numpy(w.view(np.int32)).to("cuda:0")
    out = torch.zeros((total // D, 2), dtype=torch.float32, device="cuda:0")
    ctxs = [cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32) for _ in range(world)]
    what = dict(case=case, order=order, fc=fc, D=D, world=world, calls=calls, n=n)
    try:
        for g, c in enumerate(ctxs):
            c.shard_init(g, world)
        for call in range(calls):
            for g, c in enumerate(ctxs):
                lo = (call * world + g) * n
                c.shard_begin(cb.IN_WORDS, d_w[lo:lo + n], n, out[lo // D:(lo + n) // D])
            cb.shard_exchange_local(ctxs)
            for c in ctxs:
                c.shard_finish()
        for c in ctxs:
            assert c.wait()[0] == n, what
        states = [c.get_state() for c in ctxs]
    finally:
        for c in ctxs:
            c.close()
    got = out.cpu().numpy()
    assert np.all(rel_rms(got, ref) <= TOL), what
    for b in range(1, world * calls):
        c0 = b * n // D
        seg = slice(max(c0 - 16, 0), min(c0 + 4096 // D + 16, total // D))
        if np.sqrt(np.mean(ref[seg].astype(np.float64) ** 2)) > 1e-3:
            assert np.all(rel_rms(got[seg], ref[seg]) <= TOL), (what, b)
    for st in states[1:]:
        np.testing.assert_array_equal(st, states[0])
