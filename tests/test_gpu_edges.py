"""Edge cases of the batched entry points: empty submits, submits shorter than one run (tail kernel only), many tiny
captures, the largest decimation on the shortest legal input, and a sharded call shorter than its fix-up window."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-5


def rel_rms(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.sqrt(np.mean((a - b) ** 2, axis=0)) / np.sqrt(np.mean(b ** 2, axis=0))


def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).to("cuda:0")


def test_empty_submits(cb, po, torch_cuda):
    torch = torch_cuda
    sos = po.butter_lowpass_sos(4, 4e6, 250e3)
    buf = torch.zeros(16, dtype=torch.int32, device="cuda:0")
    out = torch.zeros((16, 2), dtype=torch.float32, device="cuda:0")
    with cb.Context(sos=sos, decimation=8, parity=cb.PARITY_F32) as ctx:
        ctx.rx_submit(buf, 0, out)
        assert ctx.wait()[0] == 0
        ctx.filter_submit(cb.IN_WORDS, buf, 0, out)
        ret, res = ctx.wait()
        assert ret == 0 and res.n_filtered == 0
        np.testing.assert_array_equal(ctx.get_state(), np.zeros(8))
    with cb.Context() as ctx:
        ctx.unpack_submit(buf, nbytes=0, cs16=torch.zeros((4, 2), dtype=torch.int16, device="cuda:0"))
        assert ctx.wait()[0] == 0
    assert float(out.abs().max()) == 0.0


@pytest.mark.parametrize("order,D,n", [(4, 8, 8), (4, 8, 152), (6, 1, 7), (6, 1, 63), (6, 1, 300), (4, 32, 32), (2, 2, 66)])
def test_submits_shorter_than_a_run_and_carried_on(cb, po, torch_cuda, order, D, n):
    """n < run length: only the sequential tail kernel runs; then a long submit continues from its state."""
    torch = torch_cuda
    n2 = 1 << 16
    sos = po.butter_lowpass_sos(order, 4e6, 100e3)
    w = po.gen_words(n + n2)
    ref = po.rx_path_f32(w, sos, D)
    pad = np.zeros((n + 3) & ~3, np.uint32)
    pad[:n] = w[:n]
    o1 = torch.zeros((max(n // D, 1), 2), dtype=torch.float32, device="cuda:0")
    o2 = torch.zeros((n2 // D, 2), dtype=torch.float32, device="cuda:0")
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32) as ctx:
        ctx.filter_submit(cb.IN_WORDS, _dev(torch, pad.view(np.int32)), n, o1)
        assert ctx.wait()[0] == n
        ctx.filter_submit(cb.IN_WORDS, _dev(torch, w[n:].view(np.int32)), n2, o2)
        assert ctx.wait()[0] == n2
    got = np.concatenate([o1.cpu().numpy()[: n // D], o2.cpu().numpy()])
    # the first outputs of a filter that starts from rest are tiny: compare on the common scale of the whole stream
    scale = np.sqrt(np.mean(ref.astype(np.float64) ** 2))
    assert np.sqrt(np.mean((got - ref[: got.shape[0]]) ** 2)) / scale <= TOL


def test_a_thousand_tiny_captures(cb, po, torch_cuda):
    torch = torch_cuda
    ns, n, D = 1000, 2048, 8
    sos = po.butter_lowpass_sos(4, 4e6, 250e3)
    base = po.gen_words(n + ns)
    ws = np.stack([base[s:s + n] for s in range(ns)])            # every capture a different slice
    out = torch.zeros((ns, n // D, 2), dtype=torch.float32, device="cuda:0")
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32, n_streams=ns, chunk_bytes=4096) as ctx:
        ctx.rx_submit(_dev(torch, ws.view(np.int32)), n * 4, out, stride_bytes=n * 4, filt_stride=n // D)
        ret, res = ctx.wait()
        sr = ctx.stream_results()
    assert ret == n and all(r.n_samples == n for r in sr)
    got = out.cpu().numpy()
    for s in (0, 1, 499, 999):
        ref = po.rx_path_f32(ws[s], sos, D, native_batch_len=4096)
        assert np.all(rel_rms(got[s], ref) <= TOL), s


def test_sharded_call_shorter_than_the_fixup_window(cb, po, torch_cuda):
    """Shards of 4096 samples with the 10 kHz filter (its incoming state matters for 11 200 samples): the fix-up is
    the whole shard again, tail included, and the carry still chains."""
    torch = torch_cuda
    n, world, calls, D = 4096 + 40, 3, 3, 1
    n -= n % D
    sos = po.butter_lowpass_sos(6, 4e6, 10e3)
    total = n * world * calls
    w = po.gen_words(total)
    ref = po.rx_path_f32(w, sos, D)
    out = torch.zeros((total // D, 2), dtype=torch.float32, device="cuda:0")
    d_w = [_dev(torch, np.concatenate([w[k * n:(k + 1) * n], np.zeros(4, np.uint32)]).view(np.int32)) for k in range(world * calls)]
    ctxs = [cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32) for _ in range(world)]
    try:
        for g, c in enumerate(ctxs):
            c.shard_init(g, world)
            assert c.fixup_samples(n) == n
        for call in range(calls):
            for g, c in enumerate(ctxs):
                k = call * world + g
                c.shard_begin(cb.IN_WORDS, d_w[k], n, out[k * n // D:(k + 1) * n // D])
            cb.shard_exchange_local(ctxs)
            for c in ctxs:
                c.shard_finish()
        for c in ctxs:
            assert c.wait()[0] == n
    finally:
        for c in ctxs:
            c.close()
    scale = np.sqrt(np.mean(ref.astype(np.float64) ** 2))
    assert np.sqrt(np.mean((out.cpu().numpy() - ref) ** 2)) / scale <= TOL
