"""Time-chunk sharding of ONE long stream over the GPUs of a box (BASELINE.json configs[4]).

Rank g of G owns samples [g*n, (g+1)*n) of the call.  Unpack is pointwise; the IIR couples shards only
through the 2K-double DF-II state per rail, so:

  1. every rank filters its shard from ZERO state (fused kernel, all ranks concurrently) -> outputs + f_g,
     the zero-state end vector (reference coordinates: caribou_cuda_get_state);
  2. the hand-off: s_in(0) = state the stream carried into this call, s_in(g+1) = M s_in(g) + f_g with
     M = A^n (caribou_cuda_shard_matrix), exchanged either
       "chain":     rank g receives s_in(g) from g-1, sends s_in(g+1) to g+1  (send/recv, G-1 serial hops), or
       "allgather": one all_gather of the f_g (4K doubles each), every rank composes its own prefix;
  3. rank g re-runs the leading tiles of its shard from s_in(g) (caribou_cuda_filter_fixup): only the first
     caribou_cuda_fixup_samples() outputs can differ above 2^-60.

The collective moves 32*K bytes per rank: its cost is latency, not NVLink bandwidth.  The host-side
composition is plain numpy so that it can be tested on CPU with the gloo backend.
"""
import numpy as np


def compose_prefix(M, F, carry):
    """s_in for every shard and the state after the last one.

    M: (S, S) shard matrix A^n, per rail.  F: (G, 2, S) zero-state end vectors.  carry: (2, S) state entering
    shard 0.  Returns (s_in[G, 2, S], carry_out[2, S]).  long double accumulation, like the library's own
    end-state update."""
    M = np.asarray(M, dtype=np.longdouble)
    F = np.asarray(F, dtype=np.longdouble)
    s = np.asarray(carry, dtype=np.longdouble).copy()
    G = F.shape[0]
    s_in = np.zeros(F.shape, dtype=np.float64)
    for g in range(G):
        s_in[g] = s.astype(np.float64)
        s = s @ M.T + F[g]
    return s_in, s.astype(np.float64)


def exchange_states(f_local, M, carry, group=None, mode="allgather", device=None):
    """The hand-off step on `group` (torch.distributed; NCCL on GPUs, gloo on CPU).

    f_local: (2, S) this rank's zero-state end vector.  Returns (s_in of this rank, carry_out); carry_out is
    the stream state after the LAST shard and is identical on every rank."""
    import torch
    import torch.distributed as dist

    rank, world = dist.get_rank(group), dist.get_world_size(group)
    f_local = np.ascontiguousarray(f_local, dtype=np.float64)
    S = f_local.shape[-1]
    dev = device if device is not None else torch.device("cpu")
    if world == 1:
        s_in, out = compose_prefix(M, f_local[None], carry)
        return s_in[0], out
    if mode == "allgather":
        t = torch.from_numpy(f_local.reshape(-1)).to(dev)
        g = torch.empty(world * t.numel(), dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(g, t, group=group)
        F = g.cpu().numpy().reshape(world, 2, S)
        s_in, out = compose_prefix(M, F, carry)
        return s_in[rank], out
    if mode != "chain":
        raise ValueError("mode must be 'allgather' or 'chain'")
    # serial chain: s_in travels rank to rank; the last rank broadcasts the carry-out
    s = np.asarray(carry, dtype=np.float64).copy()
    if rank > 0:
        r = torch.empty(2 * S, dtype=torch.float64, device=dev)
        dist.recv(r, src=dist.get_global_rank(group, rank - 1) if group is not None else rank - 1, group=group)
        s = r.cpu().numpy().reshape(2, S)
    _, nxt = compose_prefix(M, f_local[None], s)
    if rank < world - 1:
        dist.send(torch.from_numpy(nxt.reshape(-1)).to(dev),
                  dst=dist.get_global_rank(group, rank + 1) if group is not None else rank + 1, group=group)
    last = torch.from_numpy(nxt.reshape(-1).copy()).to(dev)
    dist.broadcast(last, src=dist.get_global_rank(group, world - 1) if group is not None else world - 1, group=group)
    return s, last.cpu().numpy().reshape(2, S)


class ShardedStream:
    """One rank's part of a time-chunked stream.  Every call to process() takes this rank's shard of the
    next G*n samples; filter state and decimation phase carry over from call to call exactly as in the
    single-GPU Context (n must be a multiple of the decimation factor)."""

    def __init__(self, ctx, group=None, mode="allgather", device=None):
        self.ctx = ctx
        self.group = group
        self.mode = mode
        self.device = device
        self.S = 2 * ctx.n_sos
        self.carry = np.zeros((2, self.S))
        self._M = {}

    def shard_matrix(self, n):
        if n not in self._M:
            self._M[n] = self.ctx.shard_matrix(n)
        return self._M[n]

    def process(self, in_kind, x, n, out):
        """x: this rank's n input elements (device), out: n / D output elements (device)."""
        ctx = self.ctx
        ctx.set_state(np.zeros(2 * self.S))
        ctx.filter_submit(in_kind, x, n, out)
        f = ctx.get_state().reshape(2, self.S)               # waits for the kernels
        s_in, self.carry = exchange_states(f, self.shard_matrix(n), self.carry, self.group, self.mode, self.device)
        if np.any(s_in != 0.0):
            ctx.filter_fixup(in_kind, x, n, out, s_in.reshape(-1))   # leaves M s_in + f in the context
        return s_in
