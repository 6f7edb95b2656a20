"""BASELINE.json's configurations at their FULL sizes (C2: 2^26 words, C3: 2^30 CF32 samples, C4: 64 x 2^24 words, and
a 2^31-word stream for the 64-bit sample index): where the oracle would take minutes, parity is checked through
size-independent properties - oracle windows at the start and deep inside the stream (a stable filter forgets its
start), batch independence, exact power-of-two linearity, chunked == one-shot, determinism - and, for the
integer-only unpack (C2), in full against the oracle.  Inputs come from the device generator, which is itself
cross-checked against the host generator (test_gpu_compat.py)."""
import numpy as np
import pytest

from tests.test_gpu_iir import TOL, rel_rms

pytestmark = pytest.mark.gpu
SETTLE = 1 << 15          # samples after which the order-4 250 kHz (and the 25 kHz order-6) filter forgot a zero start


def _oracle_window(po, sos, n0, m, D, seed=0xC0FFEE, settle=SETTLE):
    """Oracle output for global samples [n0, n0 + m) of the generated stream, decimated by D (n0 % D == 0)."""
    lead = min(settle, n0)
    w = po.gen_words(lead + m, n0=n0 - lead, seed=seed)
    _, cs16, _, _ = po.smi_read(w.view(np.uint8), w.size, 0)
    y = po.filter_cf32_f32(sos, po.cs16_to_cf32(cs16[: w.size]))
    return y[lead:][::D]


def test_c4_batch_of_64_captures(cb, po, torch_cuda):
    """C4: 64 captures x 2^24 words, fused unpack -> order-4 LPF -> decimate by 8 in one launch."""
    torch = torch_cuda
    ns, n, D = 64, 1 << 24, 8
    sos = po.butter_lowpass_sos(4, 4e6, 250e3)
    ti, tq = po.tone_tables()
    d_w = torch.empty((ns, n), dtype=torch.int32, device="cuda:0")
    out = torch.zeros((ns, n // D, 2), dtype=torch.float32, device="cuda:0")
    torch.cuda.synchronize()
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32, n_streams=ns) as ctx:
        for c in range(ns):
            ctx.synth_words(d_w[c], n, seed=c, tone_i=ti, tone_q=tq)
        ctx.rx_submit(d_w, n * 4, out, stride_bytes=n * 4, filt_stride=n // D, mem=cb.MEM_DEVICE)
        ret, res = ctx.wait()
        assert ret == n and res.n_filtered == n // D and res.n_invalid == 0
        first = out.clone()
        torch.cuda.synchronize()                                      # torch works on its own stream
        ctx.reset()
        ctx.rx_submit(d_w, n * 4, out, stride_bytes=n * 4, filt_stride=n // D, mem=cb.MEM_DEVICE)
        ctx.wait()
    assert torch.equal(first, out)                                    # same bits every time (tile order is free)
    assert bool(torch.isfinite(out).all())
    m = 1 << 17
    for c in (0, 21, 63):
        head = out[c, : m // D].cpu().numpy()
        assert np.all(rel_rms(head, _oracle_window(po, sos, 0, m, D, seed=c)) <= TOL)
        n0 = n - m
        tail = out[c, n0 // D:].cpu().numpy()
        assert np.all(rel_rms(tail, _oracle_window(po, sos, n0, m, D, seed=c)) <= TOL)
    # batch independence: capture 63 filtered on its own gives the same bits as row 63 of the batch
    alone = torch.zeros((n // D, 2), dtype=torch.float32, device="cuda:0")
    torch.cuda.synchronize()
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32) as ctx:
        ctx.rx_submit(d_w[63], n * 4, alone, filt_stride=n // D, mem=cb.MEM_DEVICE)
        ctx.wait()
    assert torch.equal(alone, out[63])


def test_c3_cf32_stream_of_2_30(cb, po, torch_cuda):
    """C3: one CF32 stream of 2^30 samples (8 GiB), order 4, D = 8."""
    torch = torch_cuda
    n, D = 1 << 30, 8
    sos = po.butter_lowpass_sos(4, 4e6, 250e3)
    ti, tq = po.tone_tables()
    x = torch.empty((n, 2), dtype=torch.float32, device="cuda:0")
    part = 1 << 28
    d_w = torch.empty(part, dtype=torch.int32, device="cuda:0")
    with cb.Context() as gen:                                         # words -> CF32 with kernel (a), a quarter at a time
        for k in range(n // part):
            gen.synth_words(d_w, part, n0=k * part, tone_i=ti, tone_q=tq)
            gen.unpack_submit(d_w, nbytes=part * 4, cf32=x[k * part:(k + 1) * part])
            r, _ = gen.wait()
            assert r == part
    del d_w
    out = torch.zeros((n // D, 2), dtype=torch.float32, device="cuda:0")
    torch.cuda.synchronize()
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32) as ctx:
        ctx.filter_submit(cb.IN_CF32, x, n, out)
        ret, res = ctx.wait()
        assert ret == n and res.n_filtered == n // D
        m = 1 << 17
        assert np.all(rel_rms(out[: m // D].cpu().numpy(), _oracle_window(po, sos, 0, m, D)) <= TOL)
        for n0 in (n // 2 - m, n - m):
            assert np.all(rel_rms(out[n0 // D:(n0 + m) // D].cpu().numpy(), _oracle_window(po, sos, n0, m, D)) <= TOL)
        # chunked == one-shot: four calls of 2^28 with the state carried over
        ctx.reset()
        chunked = torch.zeros_like(out)
        torch.cuda.synchronize()
        for k in range(4):
            ctx.filter_submit(cb.IN_CF32, x[k * part:(k + 1) * part], part, chunked[k * part // D:(k + 1) * part // D])
            ctx.wait()
        d = (chunked - out).abs().max().item()
        assert d <= 2.0 ** -22, d                                      # |y| < 1.5: a couple of float ulps at most
        # exact linearity under a power of two: filter(4 x) == 4 filter(x), bit for bit
        ctx.reset()
        x.mul_(4.0)
        torch.cuda.synchronize()
        ctx.filter_submit(cb.IN_CF32, x, n, chunked)
        ctx.wait()
        assert torch.equal(chunked, out * 4.0)


@pytest.mark.parametrize("B", [0, 1])
def test_c2_unpack_of_2_26_words_in_full(cb, po, torch_cuda, B):
    """C2: 2^26 words -> CS16 + meta + CF32 (+ flag lists), single channel and word-interleaved dual channel,
    compared with the oracle over the whole capture."""
    torch = torch_cuda
    n = 1 << 26
    ti, tq = po.tone_tables()
    d_w = torch.empty(n, dtype=torch.int32, device="cuda:0")
    m = n if B == 0 else n // 2
    bufs = {}
    for suf in ([""] if B == 0 else ["", "_b"]):
        bufs["cs16" + suf] = torch.zeros((m, 2), dtype=torch.int16, device="cuda:0")
        bufs["meta" + suf] = torch.zeros(m, dtype=torch.uint8, device="cuda:0")
        bufs["cf32" + suf] = torch.zeros((m, 2), dtype=torch.float32, device="cuda:0")
    syn = torch.zeros(64, dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    with cb.Context(interleave_block_words=B) as ctx:
        ctx.synth_words(d_w, n, tone_i=ti, tone_q=tq)
        ctx.wait()
        ctx.unpack_submit(d_w, nbytes=n * 4, sync_idx=syn, **bufs)
        ret, res = ctx.wait()
    w = d_w.cpu().numpy().view(np.uint32)
    np.testing.assert_array_equal(w[: 1 << 20], po.gen_words(1 << 20))
    np.testing.assert_array_equal(w[-(1 << 20):], po.gen_words(1 << 20, n0=n - (1 << 20)))
    ni, _, nsy, syn_ref = po.flag_scan(w, cap=64)
    assert res.n_invalid == ni == 0 and res.n_sync == nsy == (n + 3999999) // 4000000
    np.testing.assert_array_equal(syn.cpu().numpy()[:nsy].astype(np.uint32), syn_ref)
    for sub, suf in ([(0, "")] if B == 0 else [(0, ""), (1, "_b")]):
        sw = w if B == 0 else po.demux_gather(w, B, sub)
        r, cs16, meta, _ = po.smi_read(sw.view(np.uint8), m, sub)
        assert (ret if sub == 0 else res.n_samples_b) == r == m
        np.testing.assert_array_equal(bufs["cs16" + suf].cpu().numpy(), cs16[:m])
        np.testing.assert_array_equal(bufs["meta" + suf].cpu().numpy(), meta[:m])
        np.testing.assert_array_equal(bufs["cf32" + suf].cpu().numpy(), po.cs16_to_cf32(cs16[:m]))


def test_long_stream_crosses_2_32_samples(cb, po, torch_cuda):
    """A stream longer than 32 bits of sample index (C5's shape on one GPU): 5 calls of 2^30 words regenerated in
    place, state carried; the last call's tail against the oracle generated at the 64-bit offset."""
    torch = torch_cuda
    part, D, calls = 1 << 30, 8, 5
    sos = po.butter_lowpass_sos(6, 4e6, 25e3)
    ti, tq = po.tone_tables()
    d_w = torch.empty(part, dtype=torch.int32, device="cuda:0")
    out = torch.zeros((part // D, 2), dtype=torch.float32, device="cuda:0")
    torch.cuda.synchronize()
    m = 1 << 17
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32) as ctx:
        for k in range(calls):
            ctx.synth_words(d_w, part, n0=k * part, tone_i=ti, tone_q=tq)
            ctx.rx_submit(d_w, part * 4, out, filt_stride=part // D, mem=cb.MEM_DEVICE)
            ret, _ = ctx.wait()
            assert ret == part
            if k in (1, calls - 1):      # the head of a later call depends on the carried state, the tail on the index
                head = _oracle_window(po, sos, k * part, m, D, settle=1 << 18)
                assert np.all(rel_rms(out[: m // D].cpu().numpy(), head) <= TOL)
                tail = _oracle_window(po, sos, (k + 1) * part - m, m, D, settle=1 << 18)
                assert np.all(rel_rms(out[-(m // D):].cpu().numpy(), tail) <= TOL)
