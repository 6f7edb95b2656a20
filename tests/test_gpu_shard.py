"""Time-chunk sharding with the hand-off on the device (BASELINE.json configs[4]) on ONE GPU:
`world` contexts play the ranks, caribou_cuda_shard_exchange_local stands in for the ncclAllGather, everything
else (pass 1 from zero state, device-side composition of the incoming states, fix-up from device memory, carried
state) is the code the multi-GPU path runs.  The real-NCCL run is tests/test_gpu_multi.py (>= 2 GPUs)."""
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


@pytest.mark.parametrize("order,fc,D,world", [(6, 10e3, 8, 4), (6, 10e3, 1, 3), (4, 250e3, 8, 8), (6, 50e3, 2, 2)])
def test_sharded_stream_equals_the_single_context_stream(cb, po, torch_cuda, order, fc, D, world):
    """Two calls of `world` shards each against (a) the sequential CPU oracle over the whole stream, with windows
    straddling every shard boundary looked at separately, and (b) one context fed the same stream in one piece:
    same outputs to 1e-6 of full scale, same end state."""
    torch = torch_cuda
    n, calls = 3 * 32768, 2                      # per shard and call; not a multiple of the tile size
    total = n * world * calls
    sos = po.butter_lowpass_sos(order, 4e6, fc)
    w = po.gen_words(total)
    ref = po.rx_path_f32(w, sos, D)
    d_w = _dev(torch, w.view(np.int32))
    out = torch.zeros((total // D, 2), dtype=torch.float32, device="cuda:0")
    ctxs = [cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32) for _ in range(world)]
    try:
        for g, c in enumerate(ctxs):
            c.shard_init(g, world)               # no communicator: the test moves the end vectors
        for call in range(calls):
            for g, c in enumerate(ctxs):
                lo = (call * world + g) * n
                c.shard_begin(cb.IN_WORDS, d_w[lo:lo + n], n, out[lo // D:(lo + n) // D])
            cb.shard_exchange_local(ctxs)
            for c in ctxs:
                c.shard_finish()
        for c in ctxs:
            assert c.wait()[0] == n
        states = [c.get_state() for c in ctxs]
    finally:
        for c in ctxs:
            c.close()
    got = out.cpu().numpy()
    assert np.all(rel_rms(got, ref) <= TOL)
    for b in range(1, world * calls):            # a hand-off bug hides in a whole-stream RMS
        c0 = b * n // D
        seg = slice(max(c0 - 64, 0), c0 + 4096 // D + 64)
        assert np.all(rel_rms(got[seg], ref[seg]) <= TOL), b
    one = torch.zeros((total // D, 2), dtype=torch.float32, device="cuda:0")
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32) as c1:
        c1.filter_submit(cb.IN_WORDS, d_w, total, one)
        assert c1.wait()[0] == total
        st1 = c1.get_state()
    assert float((out - one).abs().max()) <= 1e-6
    for st in states:                             # every rank carries the same stream state
        np.testing.assert_allclose(st, st1, rtol=1e-9, atol=1e-12 * np.max(np.abs(st1)))
        np.testing.assert_array_equal(st, states[0])


def test_sharded_call_through_nccl_on_one_rank(cb, po, torch_cuda):
    """caribou_cuda_shard_submit with a real communicator of one rank (ncclCommInitRank inside the library): the
    carried state makes three calls equal one long submit, CS16 input and the reference's Q16 ordering."""
    torch = torch_cuda
    n, calls, D = 1 << 17, 3, 1
    sos = po.butter_lowpass_sos(6, 4e6, 25e3)
    w = po.gen_words(n * calls)
    _, cs16, _, _ = po.smi_read(w.view(np.uint8), w.size)
    cs16 = cs16[: w.size]
    ref = po.filter_cs16_q16(sos, cs16)
    d_x = _dev(torch, cs16)
    out = torch.zeros((n * calls, 2), dtype=torch.int16, device="cuda:0")
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_Q16, out_format=cb.FMT_CS16) as c:
        c.shard_init(0, 1, cb.shard_unique_id())
        c.set_profiling(True)
        for k in range(calls):
            c.shard_submit(cb.IN_CS16, d_x[k * n:(k + 1) * n], n, out[k * n:(k + 1) * n])
        assert c.wait()[0] == n
        pass_ms, hand_ms = c.shard_profile()
        assert pass_ms > 0 and hand_ms > 0
    d = np.abs(out.cpu().numpy().astype(np.int32) - ref.astype(np.int32))
    assert d.max() <= 1 and np.mean(d > 0) < 1e-4          # int16 truncation may flip an LSB


def test_misaligned_shard_reports_sync_lost(cb, po, torch_cuda):
    """Shards consist of aligned chunks: a raw-word shard with a misaligned chunk makes the call a no-op on the
    device and wait() return -3; the carried state is left alone."""
    torch = torch_cuda
    n, D = 1 << 17, 8
    sos = po.butter_lowpass_sos(4, 4e6, 250e3)
    w = po.gen_words(2 * n)
    bad = w[n:].copy()
    bad[0] ^= 0x80000000
    out = torch.zeros((n // D, 2), dtype=torch.float32, device="cuda:0")
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32) as c:
        c.shard_init(0, 1)
        c.shard_begin(cb.IN_WORDS, _dev(torch, w[:n].view(np.int32)), n, out)
        c.shard_finish()
        assert c.wait()[0] == n
        st = c.get_state()
        c.shard_begin(cb.IN_WORDS, _dev(torch, bad.view(np.int32)), n, out)
        c.shard_finish()
        assert c.wait()[0] == cb.ERR_SYNC
        np.testing.assert_array_equal(c.get_state(), st)
