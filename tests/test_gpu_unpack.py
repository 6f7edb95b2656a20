"""Parity of kernel (a) (alignment search + unpack/demux/flags) with the CPU oracle.
Bit-exact on CS16, meta, CF32, flag counts and indices; runs through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SENT16 = 0x5A5A


def _unpack_dev(cb, torch, raw_u8, channel=0, chunk_bytes=0, B=0, cap=4096):
    """Run unpack_submit on device buffers pre-filled with sentinels. Returns dict of numpy arrays."""
    n = raw_u8.size // 4 + 2
    dev = torch.device("cuda:0")
    pad = (-raw_u8.size) % 16
    d_raw = torch.from_numpy(np.concatenate([raw_u8, np.zeros(pad + 16, np.uint8)])).to(dev)
    outs = {}
    for suf in ("", "_b"):
        outs["cs16" + suf] = torch.full((n, 2), SENT16, dtype=torch.int16, device=dev)
        outs["meta" + suf] = torch.full((n,), 0x5A, dtype=torch.uint8, device=dev)
        outs["cf32" + suf] = torch.full((n, 2), -7.0, dtype=torch.float32, device=dev)
    inv = torch.zeros(cap, dtype=torch.int32, device=dev)
    syn = torch.zeros(cap, dtype=torch.int32, device=dev)
    with cb.Context(channel=channel, chunk_bytes=chunk_bytes, interleave_block_words=B) as ctx:
        kw = dict(outs) if B else {k: v for k, v in outs.items() if not k.endswith("_b")}
        ctx.unpack_submit(d_raw, nbytes=raw_u8.size, invalid_idx=inv, sync_idx=syn, **kw)
        ret, res = ctx.wait()
    r = {k: v.cpu().numpy() for k, v in outs.items()}
    r.update(ret=ret, res=res, inv=inv.cpu().numpy().astype(np.uint32), syn=syn.cpu().numpy().astype(np.uint32))
    return r


def _oracle(po, raw_u8, channel, nbl):
    n = raw_u8.size // 4 + 2
    ret, cs16, meta, _ = po.smi_read(raw_u8, n - 1, channel, nbl, sentinel=SENT16)
    return ret, cs16[:n], meta[:n]


def _check_single(cb, po, torch, raw_u8, channel, nbl):
    g = _unpack_dev(cb, torch, raw_u8, channel, nbl)
    ret, cs16, meta = _oracle(po, raw_u8, channel, nbl)
    assert g["ret"] == ret
    np.testing.assert_array_equal(g["cs16"], cs16)
    np.testing.assert_array_equal(g["meta"], meta)
    # CF32 = cs16 / 4096 wherever the reference wrote CS16; untouched slots keep the sentinel
    written = ~((cs16[:, 0] == np.int16(SENT16)) & (cs16[:, 1] == np.int16(SENT16)))
    exp = np.where(written[:, None], po.cs16_to_cf32(cs16), np.float32(-7.0))
    np.testing.assert_array_equal(g["cf32"], exp)
    words = raw_u8[: raw_u8.size // 4 * 4].view(np.uint32)
    ni, inv, ns, syn = po.flag_scan(words, cap=4096)
    assert g["res"].n_invalid == ni and g["res"].n_sync == ns
    np.testing.assert_array_equal(g["inv"][: len(inv)], inv)
    np.testing.assert_array_equal(g["syn"][: len(syn)], syn)
    return g


@pytest.mark.parametrize("channel", [0, 1])
def test_aligned_capture(cb, po, torch_cuda, channel):
    w = po.gen_words(300000, channel=channel)
    w[::1000] |= 1          # sync bits
    _check_single(cb, po, torch_cuda, w.view(np.uint8), channel, 65536)


def test_exhaustive_codes(cb, po, torch_cuda):
    """Every 13-bit I and Q code, both S values, both layouts."""
    codes = np.arange(8192)
    for channel in (0, 1):
        w = np.concatenate([po.make_word(codes, codes[::-1], s, channel) for s in (0, 1)])
        _check_single(cb, po, torch_cuda, w.view(np.uint8), channel, 4096)


@pytest.mark.parametrize("drop", [1, 2, 3, 5, 8, 13])
def test_misaligned_stream(cb, po, torch_cuda, drop):
    w = po.gen_words(50000)
    raw = w.view(np.uint8)[drop:]
    g = _check_single(cb, po, torch_cuda, raw, 0, 16384)
    assert (g["res"].n_misaligned_chunks > 0) == (drop % 4 != 0)    # dropping whole words keeps alignment


def test_corrupt_headers_and_lost_sync(cb, po, torch_cuda):
    w = po.gen_words(40000)
    w[100] ^= 0x80000000          # mid-chunk corruption: flagged, no effect on decode
    w[4096] ^= 0x00004000         # first word of chunk 1 -> offs = 4 there
    g = _check_single(cb, po, torch_cuda, w.view(np.uint8), 0, 16384)
    assert g["res"].n_invalid == 2
    # chunk 3 entirely garbage -> -3, nothing written from there on
    w2 = po.gen_words(40000)
    w2[3 * 4096: 4 * 4096] = 0
    g = _check_single(cb, po, torch_cuda, w2.view(np.uint8), 0, 16384)
    assert g["ret"] == -3 and g["res"].first_bad_chunk == 3


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 17, 4095, 4097])
def test_ragged_lengths(cb, po, torch_cuda, n):
    w = po.gen_words(max(n, 1))[:n]
    raw = w.view(np.uint8)
    _check_single(cb, po, torch_cuda, raw, 0, 4096)
    if n >= 4:
        _check_single(cb, po, torch_cuda, raw[:-1], 0, 4096)   # byte count not a multiple of 4


@pytest.mark.parametrize("B", [1, 8, 4096, 131072])
def test_dual_channel_demux(cb, po, torch_cuda, B):
    n = 1 << 17 if B < 131072 else 1 << 20        # B = 131072: MTU-interleaved, SURVEY 8a A9's secondary case
    a = po.gen_words(n // 2, channel=0)
    b = po.gen_words(n // 2, channel=1)
    w = np.empty(n, np.uint32)
    blk = np.arange(n) // B % 2
    w[blk == 0] = a
    w[blk == 1] = b
    w[5000] ^= 0x40000000
    g = _unpack_dev(cb, torch_cuda, w.view(np.uint8), 0, 32768, B=B)
    for sub, suf in ((0, ""), (1, "_b")):
        sw = po.demux_gather(w, B, sub)
        ret, cs16, meta = _oracle(po, sw.view(np.uint8), sub, 32768)
        assert (g["ret"] if sub == 0 else g["res"].n_samples_b) == ret
        m = ret
        np.testing.assert_array_equal(g["cs16" + suf][:m], cs16[:m])
        np.testing.assert_array_equal(g["meta" + suf][:m], meta[:m])
        np.testing.assert_array_equal(g["cf32" + suf][:m], po.cs16_to_cf32(cs16[:m]))
    ni, inv, ns, syn = po.flag_scan(w, cap=4096)
    assert g["res"].n_invalid == ni == 1 and g["inv"][0] == 5000


def test_host_memory_path(cb, po, torch_cuda):
    """Same call with pageable host buffers (H2D/D2H inside the library), incl. a misaligned chunk."""
    w = po.gen_words(30000)
    raw = np.ascontiguousarray(w.view(np.uint8)[2:])
    n = raw.size // 4 + 2
    cs16 = np.full((n, 2), SENT16, np.int16)
    meta = np.full(n, 0x5A, np.uint8)
    with cb.Context(chunk_bytes=16384) as ctx:
        ctx.unpack_submit(raw, cs16=cs16, meta=meta)
        ret, _ = ctx.wait()
    eret, ecs, emeta = _oracle(po, raw, 0, 16384)
    assert ret == eret
    np.testing.assert_array_equal(cs16, ecs)
    np.testing.assert_array_equal(meta, emeta)


@pytest.mark.parametrize("drop", [0, 2, 5])
@pytest.mark.parametrize("pinned_in", [False, True])
def test_host_memory_bounce_path_leaves_untouched_slots(cb, po, torch_cuda, drop, pinned_in):
    """Pageable outputs: one synchronisation, the caller's arrays receive exactly the slots the reference writes
    (sentinels elsewhere), also with the input in page-locked memory (CARIBOU_CUDA_MEM_IO, what the shims use)."""
    torch = torch_cuda
    w = po.gen_words(100000)
    w[40000] ^= 0x80000000
    raw = np.ascontiguousarray(w.view(np.uint8)[drop:])
    n = raw.size // 4 + 2
    cs16 = np.full((n, 2), SENT16, np.int16)
    meta = np.full(n, 0x5A, np.uint8)
    cf32 = np.full((n, 2), -7.0, np.float32)
    src, mem = raw, None
    if pinned_in:
        src = torch.from_numpy(np.concatenate([raw, np.zeros((-raw.size) % 16 + 16, np.uint8)])).pin_memory()
        mem = cb.MEM_IO(cb.MEM_HOST_PINNED, cb.MEM_HOST)
    with cb.Context(chunk_bytes=16384) as ctx:
        ctx.unpack_submit(src, nbytes=raw.size, cs16=cs16, meta=meta, cf32=cf32, mem=mem)
        ret, res = ctx.wait()
    eret, ecs, emeta = _oracle(po, raw, 0, 16384)
    assert ret == eret
    np.testing.assert_array_equal(cs16, ecs)
    np.testing.assert_array_equal(meta, emeta)
    written = ~((ecs[:, 0] == np.int16(SENT16)) & (ecs[:, 1] == np.int16(SENT16)))
    np.testing.assert_array_equal(cf32, np.where(written[:, None], po.cs16_to_cf32(ecs), np.float32(-7.0)))


def test_c1_at_its_stated_size(cb, po, torch_cuda):
    """BASELINE.json configs[0]: a 4 MSPS S1G-only capture of 1 s = 2^22 words (16 MiB), unpack -> CS16 + meta + CF32,
    no filter: the whole capture against the reference's read loop, bit for bit."""
    torch = torch_cuda
    n = 1 << 22
    w = po.gen_words(n)
    d_w = torch.from_numpy(w.view(np.int32)).to("cuda:0")
    cs16 = torch.zeros((n, 2), dtype=torch.int16, device="cuda:0")
    meta = torch.zeros(n, dtype=torch.uint8, device="cuda:0")
    cf32 = torch.zeros((n, 2), dtype=torch.float32, device="cuda:0")
    with cb.Context() as ctx:
        ctx.unpack_submit(d_w, cs16=cs16, meta=meta, cf32=cf32)
        ret, res = ctx.wait()
    r_ret, r_cs16, r_meta, _ = po.smi_read(w.view(np.uint8), n)
    assert ret == r_ret == n and res.n_sync == 2 and res.n_invalid == 0      # S at n = 0 and n = 4e6
    np.testing.assert_array_equal(cs16.cpu().numpy(), r_cs16[:n])
    np.testing.assert_array_equal(meta.cpu().numpy(), r_meta[:n])
    np.testing.assert_array_equal(cf32.cpu().numpy(), po.cs16_to_cf32(r_cs16[:n]))
