"""caribou_cuda_unpack_noalloc (kernel (a) on the library's fixed scratch array, no allocation; no counterpart in
the reference) against the existing entry point caribou_cuda_unpack_submit + caribou_cuda_wait: bit-identical CS16,
meta, CF32, flag lists and result, on calls made back to back (the array keeps the previous call's contents).

Collected last on purpose.  Passes alone in a fresh process and after the rest."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SENT16 = 0x5A5A
CAP = 4096


def _buffers(torch, n):
    dev = torch.device("cuda:0")
    return dict(cs16=torch.full((n, 2), SENT16, dtype=torch.int16, device=dev),
                meta=torch.full((n,), 0x5A, dtype=torch.uint8, device=dev),
                cf32=torch.full((n, 2), -7.0, dtype=torch.float32, device=dev),
                invalid_idx=torch.full((CAP,), -1, dtype=torch.int32, device=dev),
                sync_idx=torch.full((CAP,), -1, dtype=torch.int32, device=dev))


def _existing(cb, torch, d_raw, nbytes, n, channel, chunk_bytes):
    o = _buffers(torch, n)
    with cb.Context(channel=channel, chunk_bytes=chunk_bytes) as ctx:
        ctx.unpack_submit(d_raw, nbytes=nbytes, **o)
        ret, res = ctx.wait()
    return o, ret, res


def _result_of(cb, d_res):
    r = cb.Result()
    raw = d_res.cpu().numpy().tobytes()
    assert len(raw) == C.sizeof(r)
    C.memmove(C.byref(r), raw, len(raw))
    return r


def _cases(po):
    """(raw bytes, channel, chunk_bytes): a dirty capture first - misaligned chunks, bad headers, sync bits, a ragged
    tail - then smaller clean ones, so that anything a call fails to clear shows up in the next."""
    w0 = po.gen_words(5 * 16384 + 1234, channel=0)
    w0[::997] |= 1
    w0[100:140] &= 0x7FFFFFFF                       # invalid headers inside chunk 0
    b0 = w0.view(np.uint8).copy()
    b0 = np.concatenate([b0[:2 * 65536], b0[2 * 65536 + 5:]])[:-3]     # chunk 2 onwards shifted by 5 bytes, ragged end
    w1 = po.gen_words(3 * 4096, channel=1, seed=5)
    w2 = po.gen_words(70000, channel=0, seed=9)
    w2[4 * 4096:5 * 4096] = 0                       # chunk 4 of 16 KiB chunks has no alignment at all: -3
    w3 = po.gen_words(1 << 20, channel=0, seed=11)
    w3[::4000] |= 1
    return [(b0, 0, 65536), (w1.view(np.uint8), 1, 16384), (w2.view(np.uint8), 0, 16384), (w3.view(np.uint8), 0, 0)]


def test_noalloc_entry_point_equals_unpack_submit_on_calls_in_a_row(cb, po, torch_cuda):
    torch = torch_cuda
    dev = torch.device("cuda:0")
    side = torch.cuda.Stream(device=dev)            # the caller's stream (second half of the calls)
    cases = _cases(po)
    prepared = []
    for raw, channel, chunk in cases:
        n = raw.size // 4 + 2
        d_raw = torch.from_numpy(np.concatenate([raw, np.zeros((-raw.size) % 16 + 16, np.uint8)])).to(dev)
        want, ret, res = _existing(cb, torch, d_raw, raw.size, n, channel, chunk)
        prepared.append((d_raw, raw.size, n, channel, chunk, want, ret, res, _buffers(torch, n),
                         torch.zeros(C.sizeof(cb.Result), dtype=torch.uint8, device=dev)))
    torch.cuda.synchronize()
    assert prepared[0][7].n_misaligned_chunks > 0 and prepared[0][7].n_invalid > 0 and prepared[2][6] == cb.ERR_SYNC
    # two rounds over all cases, back to back, nothing waited for in between: 8 calls in a row on the shared array
    for rnd in range(2):
        for k, (d_raw, nbytes, n, channel, chunk, want, ret, res, got, d_res) in enumerate(prepared):
            if rnd == 1:
                fresh = _buffers(torch, n)
                for name in got:
                    got[name].copy_(fresh[name])
                d_res.zero_()
                side.wait_stream(torch.cuda.current_stream())
            # round 0 on the default stream (NULL), round 1 on the caller's own stream
            cb.unpack_noalloc(d_raw, nbytes, result=d_res, stream=side.cuda_stream if rnd == 1 else None,
                              channel=channel, chunk_bytes=chunk, **got)
        side.synchronize()
        torch.cuda.synchronize()
        for k, (d_raw, nbytes, n, channel, chunk, want, ret, res, got, d_res) in enumerate(prepared):
            for name in want:
                assert torch.equal(got[name], want[name]), (rnd, k, name)
            r = _result_of(cb, d_res)
            assert (r.n_samples, r.n_samples_b, r.n_filtered) == (ret, 0, 0), (rnd, k)
            assert (r.n_invalid, r.n_sync) == (res.n_invalid, res.n_sync), (rnd, k)
            assert (r.first_bad_chunk, r.n_misaligned_chunks) == (res.first_bad_chunk, res.n_misaligned_chunks), (rnd, k)


def test_noalloc_entry_point_equals_the_oracle(cb, po, torch_cuda):
    """Independent of the other entry point: the reference's read (caribou_smi.c:632-682) on the same bytes."""
    torch = torch_cuda
    dev = torch.device("cuda:0")
    for raw, channel, chunk in _cases(po)[:3]:
        n = raw.size // 4 + 2
        d_raw = torch.from_numpy(np.concatenate([raw, np.zeros((-raw.size) % 16 + 16, np.uint8)])).to(dev)
        got = _buffers(torch, n)
        d_res = torch.zeros(C.sizeof(cb.Result), dtype=torch.uint8, device=dev)
        cb.unpack_noalloc(d_raw, raw.size, result=d_res, channel=channel, chunk_bytes=chunk, **got)
        torch.cuda.synchronize()
        ret, cs16, meta, _ = po.smi_read(raw, n - 1, channel, chunk, sentinel=SENT16)
        assert _result_of(cb, d_res).n_samples == ret
        np.testing.assert_array_equal(got["cs16"].cpu().numpy(), cs16[:n])
        np.testing.assert_array_equal(got["meta"].cpu().numpy(), meta[:n])


def test_noalloc_entry_point_takes_no_device_memory_and_rejects_oversize(cb, po, torch_cuda):
    torch = torch_cuda
    dev = torch.device("cuda:0")
    w = po.gen_words(1 << 18)
    d_raw = torch.from_numpy(w.view(np.int32)).to(dev)
    got = _buffers(torch, w.size + 2)
    cb.unpack_noalloc(d_raw, w.nbytes, **got)                  # first use on this device: module load, symbol lookup
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    for _ in range(3):
        cb.unpack_noalloc(d_raw, w.nbytes, **got)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 >= free0                                      # nothing was allocated on the device
    # beyond the fixed scratch: a clear error, nothing launched, outputs untouched
    before = got["cs16"].clone()
    got["cs16"].fill_(123)
    with pytest.raises(cb.CaribouCudaError, match="exceed the fixed scratch"):
        cb.unpack_noalloc(d_raw, cb.NOALLOC_MAX_BYTES + 16, **got)
    with pytest.raises(cb.CaribouCudaError, match="chunks exceed the fixed scratch"):
        cb.unpack_noalloc(d_raw, w.nbytes, chunk_bytes=32, **got)          # 32768 chunks > 16384
    torch.cuda.synchronize()
    assert bool((got["cs16"] == 123).all()) and before.shape == got["cs16"].shape
    # the largest size the array covers goes through (2^24 words, default chunks)
    wb = po.gen_words(cb.NOALLOC_MAX_BYTES // 4, seed=21)
    wb[::4000000] |= 1
    big = torch.from_numpy(wb.view(np.int32)).to(dev)
    o_big = _buffers(torch, big.numel())
    d_res = torch.zeros(C.sizeof(cb.Result), dtype=torch.uint8, device=dev)
    cb.unpack_noalloc(big, cb.NOALLOC_MAX_BYTES, result=d_res, cs16=o_big["cs16"], meta=o_big["meta"], cf32=o_big["cf32"])
    want, ret, res = _existing(cb, torch, big, cb.NOALLOC_MAX_BYTES, big.numel(), 0, 0)
    torch.cuda.synchronize()
    r = _result_of(cb, d_res)
    assert r.n_samples == ret and (r.n_invalid, r.n_sync) == (res.n_invalid, res.n_sync)
    for name in ("cs16", "meta", "cf32"):
        assert torch.equal(o_big[name], want[name]), name


@pytest.mark.parametrize("case", list(range(16)))
def test_random_noalloc_unpack(cb, po, torch_cuda, case):
    """caribou_cuda_unpack_noalloc against caribou_cuda_unpack_submit on the same damaged capture (random channel, chunk
    size, length up to a few MiB, dropped bytes, corrupted headers), back to back on the shared scratch array."""
    import ctypes as C
    torch = torch_cuda
    rng = np.random.default_rng(33000 + case + 100000 * int(__import__("os").environ.get("CARIBOU_FUZZ_SEED", "0")))
    dev = torch.device("cuda:0")
    for rep in range(3):
        channel = int(rng.integers(2))
        chunk = int([256, 4096, 65536, 0][rng.integers(4)])
        n = int(rng.integers(1, 400000 if chunk != 256 else 300000))
        w = po.gen_words(n, channel=channel, seed=case * 7 + rep)
        w[rng.integers(0, n, size=max(1, n // 800))] |= 1
        for _ in range(int(rng.integers(0, 3))):
            w[int(rng.integers(n))] ^= 0x80000000
        raw = w.view(np.uint8).copy()
        if rng.integers(2):
            p = int(rng.integers(raw.size))
            raw = np.concatenate([raw[:p], raw[p + int(rng.integers(1, 4)):]])
        if (raw.size + max(chunk, 1) - 1) // max(chunk or 524288, 1) > cb.NOALLOC_MAX_CHUNKS:
            continue
        m = raw.size // 4 + 2
        d_raw = torch.from_numpy(np.concatenate([raw, np.zeros((-raw.size) % 16 + 16, np.uint8)])).to(dev)

        def bufs():
            return dict(cs16=torch.full((m, 2), 0x5A5A, dtype=torch.int16, device=dev),
                        meta=torch.full((m,), 0x5A, dtype=torch.uint8, device=dev),
                        cf32=torch.full((m, 2), -7.0, dtype=torch.float32, device=dev),
                        invalid_idx=torch.full((1024,), -1, dtype=torch.int32, device=dev),
                        sync_idx=torch.full((1024,), -1, dtype=torch.int32, device=dev))
        want = bufs()
        with cb.Context(channel=channel, chunk_bytes=chunk) as ctx:
            ctx.unpack_submit(d_raw, nbytes=raw.size, **want)
            ret, res = ctx.wait()
        got = bufs()
        d_res = torch.zeros(C.sizeof(cb.Result), dtype=torch.uint8, device=dev)
        cb.unpack_noalloc(d_raw, raw.size, result=d_res, channel=channel, chunk_bytes=chunk, **got)
        torch.cuda.synchronize()
        r = cb.Result()
        C.memmove(C.byref(r), d_res.cpu().numpy().tobytes(), C.sizeof(r))
        what = dict(case=case, rep=rep, channel=channel, chunk=chunk, n=n)
        assert (r.n_samples, r.n_invalid, r.n_sync, r.first_bad_chunk, r.n_misaligned_chunks) == \
               (ret, res.n_invalid, res.n_sync, res.first_bad_chunk, res.n_misaligned_chunks), what
        for name in want:
            assert torch.equal(got[name], want[name]), (what, name)
