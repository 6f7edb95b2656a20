#!/usr/bin/env python
"""PCIe ceiling of the box, to put the host-buffer (`e2e`) figure in context: pinned H2D of 4 GiB, pinned D2H of
1 GiB, and both at once on two streams (the shape of one C4 step).  One JSON line."""
import json
import torch

dev = torch.device("cuda", 0)
h_in = torch.empty(4 << 30, dtype=torch.uint8).pin_memory()
h_out = torch.empty(1 << 30, dtype=torch.uint8).pin_memory()
d_in = torch.empty(4 << 30, dtype=torch.uint8, device=dev)
d_out = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)


def timed(fn, reps=5):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        for s in (s1, s2):
            torch.cuda.current_stream().wait_stream(s)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def h2d():
    s1.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def d2h():
    s2.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def both():
    h2d(); d2h()


t_h2d, t_d2h, t_both = timed(h2d), timed(d2h), timed(both)
print(json.dumps({"h2d_4GiB_ms": t_h2d, "h2d_GBps": (4 << 30) / t_h2d / 1e6, "d2h_1GiB_ms": t_d2h,
                  "d2h_GBps": (1 << 30) / t_d2h / 1e6, "both_ms": t_both,
                  "c4_step_ceiling_Msps": (1 << 30) / t_both / 1e3}))
