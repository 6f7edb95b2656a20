"""The box schedule of the rerun kernel's ring (cariboulite_b200/csrc/iir_rerun.cuh), checked as a model on the CPU.

A warp owns a ring of R shared-memory slots fed by TMA; a tile is N boxes.  Phase 1 consumes the boxes in the order
keep .. N-1, 0 .. keep-1, phase 3 in the order 0 .. N-1; the `keep` boxes phase 1 looks at last stay in the ring for
phase 3, the others are requested a second time; a slot whose box is done with goes to the next request of the tile's
sequence and then to the next tile's.  A gated tile has no phase 3.  Each slot has an mbarrier whose phase completes
once per request; the kernel keeps one parity bit per slot, flipped when the slot is handed to a new request.

The model below is the kernel's bookkeeping, statement by statement (slot pointer, parity mask, request index `li`,
seq_box, rerun_keep); what it asserts at every consumption is what the kernel relies on: the slot holds the box the
phase wants, and the parity bit names the barrier phase of exactly that request."""
import itertools

import pytest


def rerun_keep(boxes, ring, ap, resident=True):
    """csrc/iir.cuh rerun_keep()."""
    if not resident:
        return 0
    return boxes if boxes <= ring else (0 if ap else ring)


def run_warp(N, ring, ap, live_flags, resident=True):
    R = min(N, ring)
    keep = rerun_keep(N, ring, ap, resident)
    LEN = 2 * N - keep

    def seq_box(idx):
        if idx < N:
            return idx + keep if idx + keep < N else idx + keep - N
        return keep + (idx - N)

    # the hardware side: what each slot holds and how many requests into it have completed
    held = [None] * R            # (tile, box)
    completed = [0] * R
    requests = []                # every request made, for the traffic count

    def issue(tile, box, slot):
        held[slot] = (tile, box)
        completed[slot] += 1     # (the model completes a request at once; the kernel waits on the barrier)
        requests.append((tile, box))

    def consume(slot, pm, want):
        assert held[slot] == want, (held[slot], want)
        # the barrier phase the kernel waits for is the one the latest request into the slot completed
        assert (pm >> slot) & 1 == (completed[slot] - 1) & 1

    slot, pm = 0, 0
    n_tiles = len(live_flags)
    tile = 0
    for s in range(R):
        issue(tile, seq_box(s), s)
    while tile < n_tiles:
        live = live_flags[tile]
        nxt = tile + 1
        # ---- phase 1
        li = R
        for j in range(N):
            st = slot
            consume(st, pm, (tile, seq_box(j)))
            frees = (not live) or j < N - keep
            if frees and li < (LEN if live else N):
                issue(tile, seq_box(li), st)
                li += 1
                pm ^= 1 << st
            slot = (slot + 1) % R
        if not live:
            if nxt < n_tiles:
                for s in range(R):
                    issue(nxt, seq_box(s), (slot + s) % R)
            pm ^= (1 << R) - 1
            tile = nxt
            continue
        assert li == R + N - keep
        # ---- phase 3
        for b in range(N):
            st = slot
            consume(st, pm, (tile, b))
            if li < LEN:
                issue(tile, seq_box(li), st)
            elif nxt < n_tiles:
                issue(nxt, seq_box(li - LEN), st)
            li += 1
            pm ^= 1 << st
            slot = (slot + 1) % R
        assert li == LEN + R
        tile = nxt
    return requests, keep


@pytest.mark.parametrize("ring", [3, 4])
@pytest.mark.parametrize("ap", [False, True])
def test_every_consumption_finds_its_box_and_its_barrier_phase(ring, ap):
    for N in range(1, 18):
        for live_flags in ([True] * 5, [True, False, True, True, False], [False, False, True], [True, True, False]):
            run_warp(N, ring, ap, live_flags)


def test_request_counts():
    """Live tiles request 2N - keep boxes (keep = N: a tile that fits the ring is fetched once), gated tiles N."""
    for N, ring, ap in itertools.product(range(1, 18), (3, 4), (False, True)):
        flags = [True, True, False, True]
        req, keep = run_warp(N, ring, ap, flags)
        per_tile = [sum(1 for t, _ in req if t == k) for k in range(len(flags))]
        assert per_tile == [2 * N - keep if f else N for f in flags], (N, ring, ap, per_tile)
        assert keep == (N if N <= ring else (0 if ap else ring))
    # without residency (CBL_RR_RESIDENT = 0): the plain order, every box twice
    req, keep = run_warp(8, 4, False, [True, True], resident=False)
    assert keep == 0 and len(req) == 32
