"""Host side of the multi-GPU time-chunk sharding (cariboulite_b200/shard.py) on CPU: two gloo ranks.

Every rank plays its GPU with the oracle cascade: shard filtered from zero state -> f_g, hand-off
(all_gather or send/recv chain), shard re-run from the received state.  The stitched result must equal the
single sequential oracle run bit for bit when the re-run covers the whole shard (it is the same arithmetic
from the same state), and the state the ranks agree on must equal the oracle's end state to rounding."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _shard_matrix(po, sos, n):
    """A^n in DF-II (v1, v2 per stage) coordinates, column by column from the oracle itself: the end state of
    n zero-input samples started from a unit state."""
    S = 2 * sos.shape[0]
    M = np.zeros((S, S))
    for j in range(S):
        c = po.Cascade(sos)
        e = np.zeros(S)
        e[j] = 1.0
        c.state = e
        c.run_f64(np.zeros(n))
        M[:, j] = c.state
    return M


def _worker(rank, world, port, mode, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from oracle import pyoracle as po
    from cariboulite_b200 import shard

    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        n, calls = 1 << 12, 3                      # per-rank shard per call; short shards: M is far from zero
        sos = po.butter_lowpass_sos(6, 4e6, 10e3)
        S = 2 * sos.shape[0]
        rng = np.random.default_rng(7)
        x = rng.integers(-4096, 4096, size=(calls * world * n, 2)).astype(np.float64)
        M = _shard_matrix(po, sos, n)
        carry = np.zeros((2, S))
        outs = []
        for call in range(calls):
            lo = (call * world + rank) * n
            xs = x[lo:lo + n]
            # 1. zero-state pass
            f = np.zeros((2, S))
            for r in range(2):
                c = po.Cascade(sos)
                c.run_f64(np.ascontiguousarray(xs[:, r]))
                f[r] = c.state
            # 2. hand-off
            s_in, carry = shard.exchange_states(f, M, carry, None, mode)
            # 3. re-run from the true incoming state
            y = np.zeros((n, 2))
            for r in range(2):
                c = po.Cascade(sos)
                c.state = s_in[r]
                y[:, r] = c.run_f64(np.ascontiguousarray(xs[:, r]))
            outs.append((lo, y))
        q.put((rank, outs, carry))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["allgather", "chain"])
def test_two_rank_handoff_matches_sequential_oracle(po, mode):
    import torch.multiprocessing as mp

    world, n, calls = 2, 1 << 12, 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, mode, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0

    sos = po.butter_lowpass_sos(6, 4e6, 10e3)
    rng = np.random.default_rng(7)
    x = rng.integers(-4096, 4096, size=(calls * world * n, 2)).astype(np.float64)
    ref = np.zeros_like(x)
    end = np.zeros((2, 2 * sos.shape[0]))
    for r in range(2):
        c = po.Cascade(sos)
        ref[:, r] = c.run_f64(np.ascontiguousarray(x[:, r]))
        end[r] = c.state
    got = np.zeros_like(x)
    for rank, outs, carry in res:
        for lo, y in outs:
            got[lo:lo + n] = y
        # the carried state is M s + f composed in long double: equal to the sequential end state to rounding
        assert np.allclose(carry, end, rtol=1e-9, atol=1e-9 * np.abs(end).max()), (rank, carry, end)
    err = np.sqrt(np.mean((got - ref) ** 2, axis=0)) / np.sqrt(np.mean(ref ** 2, axis=0))
    assert np.all(err < 1e-12), err


def test_compose_prefix_is_the_affine_chain():
    from cariboulite_b200 import shard
    rng = np.random.default_rng(1)
    S, G = 6, 5
    M = rng.normal(size=(S, S)) * 0.3
    F = rng.normal(size=(G, 2, S))
    carry = rng.normal(size=(2, S))
    s_in, out = shard.compose_prefix(M, F, carry)
    s = carry.copy()
    for g in range(G):
        assert np.allclose(s_in[g], s)
        s = np.stack([M @ s[0] + F[g, 0], M @ s[1] + F[g, 1]])
    assert np.allclose(out, s)
