"""Continuous-streaming queue without a GPU: caribou_cuda_ring against the reference's own
circular_buffer<T> (LC/datatypes/circular_buffer.h, compiled unmodified into oracle/_ref) on random operation
sequences in all four (override_write, block_read) modes, against a plain Python model of it, and across threads."""
import ctypes as C
import threading

import numpy as np
import pytest


class Model:
    """circular_buffer.h:16-92 as a list."""

    def __init__(self, size, override, block):
        cap = 1
        while cap < size:
            cap <<= 1
        self.cap, self.override, self.block, self.q = cap, override, block, []

    def put(self, data):
        free = self.cap - len(self.q)
        if free < len(data) and self.override:
            del self.q[: len(data) - free]
        n = min(len(data), self.cap - len(self.q))
        self.q.extend(data[:n])
        return n

    def get(self, length):
        if self.block and len(self.q) < length:
            return []
        n = min(length, len(self.q))
        out, self.q = self.q[:n], self.q[n:]
        return out


@pytest.mark.parametrize("override", [1, 0])
@pytest.mark.parametrize("block", [1, 0])
@pytest.mark.parametrize("size", [8, 100, 1024])
def test_ring_matches_reference_circular_buffer(cb, po, override, block, size):
    L = cb.lib()
    R = po.ref() if po.ref_available() else None
    mine = L.caribou_cuda_ring_create(size, 4, override, block)
    ref = R.ref_ring_create(size, override, block) if R else None
    model = Model(size, override, block)
    cap = L.caribou_cuda_ring_capacity(mine)
    assert cap == model.cap and (not R or R.ref_ring_capacity(ref) == cap)
    rng = np.random.default_rng(size * 4 + override * 2 + block)
    counter = 0
    try:
        for step in range(3000):
            op = rng.random()
            if op < 0.5:
                n = int(rng.integers(0, cap + 1))                    # the reference overruns its buffer beyond cap
                data = np.arange(counter, counter + n, dtype=np.uint32)
                counter += n
                exp = model.put(list(data))
                assert L.caribou_cuda_ring_put(mine, data.ctypes.data, n) == exp
                if R:
                    assert R.ref_ring_put(ref, data.ctypes.data, n) == exp
            elif op < 0.97:
                n = int(rng.integers(0, cap + 2))
                exp = model.get(n)
                a = np.full(n + 1, 0xDEADBEEF, dtype=np.uint32)
                null = rng.random() < 0.1                            # data == NULL: consume only
                got = L.caribou_cuda_ring_get(mine, None if null else a.ctypes.data, n, 200)
                assert got == len(exp)
                if not null:
                    np.testing.assert_array_equal(a[:got], np.array(exp, dtype=np.uint32))
                    assert a[got] == 0xDEADBEEF
                if R:
                    b = np.full(n + 1, 0xDEADBEEF, dtype=np.uint32)
                    assert R.ref_ring_get(ref, None if null else b.ctypes.data, n, 200) == got
                    if not null:
                        np.testing.assert_array_equal(b[:got], a[:got])
            else:
                model.q = []
                L.caribou_cuda_ring_reset(mine)
                if R:
                    R.ref_ring_reset(ref)
            assert L.caribou_cuda_ring_size(mine) == len(model.q)
            if R:
                assert R.ref_ring_size(ref) == len(model.q)
    finally:
        L.caribou_cuda_ring_destroy(mine)
        if R:
            R.ref_ring_destroy(ref)


def test_ring_put_longer_than_capacity_keeps_the_newest(cb):
    L = cb.lib()
    r = L.caribou_cuda_ring_create(16, 4, 1, 0)
    data = np.arange(50, dtype=np.uint32)
    assert L.caribou_cuda_ring_put(r, data.ctypes.data, 50) == 16
    out = np.zeros(16, dtype=np.uint32)
    assert L.caribou_cuda_ring_get(r, out.ctypes.data, 16, 0) == 16
    np.testing.assert_array_equal(out, data[-16:])
    assert L.caribou_cuda_ring_dropped(r) == 34
    L.caribou_cuda_ring_destroy(r)


def test_ring_blocking_get_across_threads(cb):
    """A producer thread and a blocking consumer: every element arrives once, in order; a get that cannot be
    satisfied within its timeout takes nothing."""
    L = cb.lib()
    r = L.caribou_cuda_ring_create(1 << 14, 4, 1, 1)
    total, blk = 200000, 1000

    def producer():
        for k in range(0, total, blk):
            d = np.arange(k, k + blk, dtype=np.uint32)
            while L.caribou_cuda_ring_size(r) + blk > (1 << 14):     # never overrun: nothing may be dropped here
                pass
            assert L.caribou_cuda_ring_put(r, d.ctypes.data, blk) == blk

    t = threading.Thread(target=producer)
    t.start()
    got = []
    buf = np.zeros(777, dtype=np.uint32)
    while sum(len(g) for g in got) < total - 777:
        n = L.caribou_cuda_ring_get(r, buf.ctypes.data, 777, 2000000)
        assert n == 777
        got.append(buf[:n].copy())
    t.join()
    rest = total - sum(len(g) for g in got)
    assert L.caribou_cuda_ring_get(r, buf.ctypes.data, rest + 1, 1000) == 0       # one more than there is: nothing taken
    assert L.caribou_cuda_ring_get(r, buf.ctypes.data, rest, 1000) == rest
    got.append(buf[:rest].copy())
    np.testing.assert_array_equal(np.concatenate(got), np.arange(total, dtype=np.uint32))
    assert L.caribou_cuda_ring_dropped(r) == 0
    L.caribou_cuda_ring_destroy(r)
