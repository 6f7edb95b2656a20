"""TX direction on the GPU: caribou_cuda_tx_pack_submit against the oracle bit for bit, the
caribou_smi_write / Stream::WriteSamplesGen shims against the oracle and against the reference's committed
behaviour (tests/golden/tx_golden.npz)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "tx_golden.npz")
NP_DT = {"CS16": np.int16, "CS8": np.int8, "CF32": np.float32, "CF64": np.float64}


def make_samples(fmt, n, seed=1):
    rng = np.random.default_rng(seed)
    if fmt == "CS16":
        x = rng.integers(-32768, 32768, size=(n, 2)).astype(np.int16)      # beyond 13 bits too: the pack truncates
        x[: min(n, 8)] = np.array([[-1, 0], [0, -1], [4095, -4096], [1, 1], [-4096, 4095], [0, 0], [8191, -8192],
                                   [32767, -32768]], dtype=np.int16)[: min(n, 8)]
        return x
    if fmt == "CS8":
        return rng.integers(-128, 128, size=(n, 2)).astype(np.int8)
    x = rng.random((n, 2)) * 2.4 - 1.2                                     # a little beyond the +-1 full scale
    if n >= 4:
        x[:4] = [[1.0, -1.0], [0.99999, -0.99999], [7.9, -8.0], [1e9, float("nan")]]
    return x.astype(NP_DT[fmt])


def expect(po, fmt, x):
    return po.tx_pack(x if fmt == "CS16" else po.tx_to_cs16(x, fmt))


@pytest.mark.parametrize("fmt", ["CS16", "CS8", "CF32", "CF64"])
@pytest.mark.parametrize("n", [1, 3, 4, 5, 4099, (1 << 20) + 1])
def test_tx_pack_host_buffers(cb, po, torch_cuda, fmt, n):
    x = make_samples(fmt, n)
    out = np.zeros(n, dtype=np.uint32)
    with cb.Context() as ctx:
        ctx.tx_pack_submit(getattr(cb, "FMT_" + fmt), x, n, out)
        ret, _ = ctx.wait()
        assert ret == n
        assert ctx.launch_count == 1
    np.testing.assert_array_equal(out, expect(po, fmt, x))


@pytest.mark.parametrize("fmt", ["CS16", "CF32"])
def test_tx_pack_device_buffers(cb, po, torch_cuda, fmt):
    torch = torch_cuda
    n = (1 << 22) + 2
    x = make_samples(fmt, n, seed=2)
    d_in = torch.from_numpy(x).cuda()
    d_out = torch.zeros(n, dtype=torch.int32, device="cuda:0")
    with cb.Context() as ctx:
        ctx.tx_pack_submit(getattr(cb, "FMT_" + fmt), d_in, n, d_out)
        ctx.wait()
        with pytest.raises(cb.CaribouCudaError):                       # misaligned device pointers are refused
            ctx.tx_pack_submit(getattr(cb, "FMT_" + fmt), d_in, n - 1, d_out[1:])
    np.testing.assert_array_equal(d_out.cpu().numpy().view(np.uint32), expect(po, fmt, x))


def _read_all(fd):
    size = os.lseek(fd, 0, os.SEEK_END)
    os.lseek(fd, 0, os.SEEK_SET)
    return np.frombuffer(os.read(fd, size), dtype=np.uint32).copy()


def test_smi_write_shim(cb, po, torch_cuda):
    """int caribou_smi_write(...) (caribou_smi.c:720-762): chunked by native_batch_len, returns samples written;
    on the reference's own test pattern (every sample (-1, 0)) the bytes equal what the reference writes."""
    L = cb.lib()
    g = np.load(GOLD)
    nbl = int(g["nbl"])
    fd = os.memfd_create("smi_tx")
    h = L.caribou_cuda_smi_open(fd, nbl, 0)
    assert h, cb.last_error()
    try:
        for n in (1, 5, 8192, 8192 * 3 + 77):
            os.ftruncate(fd, 0); os.lseek(fd, 0, os.SEEK_SET)
            x = np.tile(np.array([[-1, 0]], dtype=np.int16), (n, 1))
            assert L.caribou_cuda_smi_write(h, 0, x.ctypes.data, n) == int(g["ret_%d" % n])
            np.testing.assert_array_equal(_read_all(fd), g["words_%d" % n])
        os.ftruncate(fd, 0); os.lseek(fd, 0, os.SEEK_SET)
        n = 8192 * 5 + 3
        x = make_samples("CS16", n, seed=3)
        assert L.caribou_cuda_smi_write(h, 1, x.ctypes.data, n) == n
        np.testing.assert_array_equal(_read_all(fd), po.tx_pack(x))
        assert L.caribou_cuda_smi_write(h, 0, None, 0) == 0
    finally:
        L.caribou_cuda_smi_close(h)
        os.close(fd)


@pytest.mark.parametrize("fmt", ["CS16", "CS8", "CF32", "CF64"])
def test_txstream_write_shim(cb, po, torch_cuda, fmt):
    """Stream::WriteSamplesGen (CaribouliteStream.cpp:199-258): the converted formats are clamped to the MTU, CS16
    is not; an unknown format writes nothing."""
    L = cb.lib()
    nbl = 32768
    mtu = nbl // 4
    fd = os.memfd_create("smi_tx")
    h = L.caribou_cuda_smi_open(fd, nbl, 0)
    assert h, cb.last_error()
    try:
        n = mtu * 2 + 100
        x = make_samples(fmt, n, seed=4)
        r = L.caribou_cuda_txstream_write(h, 0, fmt.encode(), x.ctypes.data, n)
        exp_n = n if fmt == "CS16" else mtu
        assert r == exp_n
        np.testing.assert_array_equal(_read_all(fd), expect(po, fmt, x[:exp_n]))
        assert L.caribou_cuda_txstream_write(h, 0, b"XX", x.ctypes.data, n) == 0
    finally:
        L.caribou_cuda_smi_close(h)
        os.close(fd)
