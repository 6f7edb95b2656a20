"""The order of the sharded hand-off in peer mode (cariboulite_b200/csrc/shard.cu), checked as a model on the CPU.

Every rank queues, on its one stream, for call k:  pass 1 of call k, composition (+ fix-up) of call k - 1, push of
call k.  A push writes this rank's end vector of call k into slot k & 1 of EVERY rank's exchange buffer and raises
that rank's flag to k; a composition of call k waits until all flags of slot k & 1 say k and then reads the slot.
The buffers are two deep, so the model has to show, for every interleaving of the ranks' streams, that
  * a composition never reads a slot a later push has already overwritten, and
  * no rank ever waits on a flag that cannot come (no deadlock),
which is the argument in the comment of caribou_cuda_shard_submit.  The alternative order (push before the pending
composition) is shown to break: a fast rank overwrites slot and flag of a call a slow rank has not composed yet, which
then waits for a flag value that is gone (the kernel would give up after its bounded spin and report the peer error)."""
import random

import pytest


def simulate(world, calls, rng, push_first=False):
    # per-rank program: list of ops in stream order
    prog = []
    for g in range(world):
        ops = []
        for k in range(1, calls + 1):
            ops.append(("pass1", k))
            late = [("compose", k - 1)] if k > 1 else []
            ops += ([("push", k)] + late) if push_first else (late + [("push", k)])
        ops.append(("compose", calls))            # the flush at caribou_cuda_wait
        prog.append(ops)
    pc = [0] * world
    slot = [[[None] * world for _ in range(2)] for _ in range(world)]      # slot[rank][parity][src] = call number held
    flag = [[[0] * world for _ in range(2)] for _ in range(world)]
    violations = 0
    while any(pc[g] < len(prog[g]) for g in range(world)):
        ready = []
        for g in range(world):
            if pc[g] >= len(prog[g]):
                continue
            op, k = prog[g][pc[g]]
            if op == "compose" and not all(flag[g][k & 1][j] == k for j in range(world)):
                continue                          # the kernel spins on the flags
            ready.append(g)
        if not ready:
            return -1                             # every unfinished rank waits for a flag that cannot come
        g = rng.choice(ready)
        op, k = prog[g][pc[g]]
        if op == "push":
            for r in range(world):
                slot[r][k & 1][g] = k
                flag[r][k & 1][g] = k
        elif op == "compose":
            if any(slot[g][k & 1][j] != k for j in range(world)):
                violations += 1                   # a newer push got there first
        pc[g] += 1
    return violations


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_composition_always_reads_its_own_call(world):
    rng = random.Random(world)
    for trial in range(300):
        assert simulate(world, calls=rng.randint(1, 7), rng=rng) == 0


def test_pushing_before_the_pending_composition_would_break():
    """With push(k) queued before composition(k - 1) a fast rank overwrites a slot a slow rank has not read yet."""
    rng = random.Random(7)
    outcomes = [simulate(3, calls=6, rng=rng, push_first=True) for _ in range(300)]
    assert any(o != 0 for o in outcomes)
