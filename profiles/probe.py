#!/usr/bin/env python
"""Kernel-time probe used while tuning (not the bench): device-resident timings of the hot-path kernels
for a handful of shapes, through the C ABI, CUDA events on the launching stream.  One JSON line per case.

    python profiles/probe.py [case ...]     cases: fused:<captures>:<log2 words>:<order>:<D>  cf32:<log2 n>:<order>:<D>
                                                   unpack:<log2 words>:<B>  tx:<log2 n>:<CS16|CS8|CF32|CF64>
"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import cariboulite_b200 as cb

PEAK = 6447.5
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
FC = {2: 250e3, 4: 250e3, 6: 25e3, 8: 400e3}


def tone_tables():
    n = np.arange(800, dtype=np.float64)
    # same shape as the oracle generator; exact values do not matter for timing
    ti = np.round(3000 * np.cos(2 * np.pi * 5e3 * n / 4e6) + 800 * np.cos(2 * np.pi * 1e6 * n / 4e6)).astype(np.int16)
    tq = np.round(3000 * np.sin(2 * np.pi * 5e3 * n / 4e6) + 800 * np.sin(2 * np.pi * 1e6 * n / 4e6)).astype(np.int16)
    return ti, tq


def timeit(stream, fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); fn(); e1.record(stream)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(np.min(ts))


def main():
    cases = sys.argv[1:] or ["fused:64:24:4:8"]
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    ti, tq = tone_tables()
    for case in cases:
        f = case.split(":")
        if f[0] == "fused":
            ns, lg, order, D = int(f[1]), int(f[2]), int(f[3]), int(f[4])
            n = 1 << lg
            sos = cb.butter_lowpass(order, 4e6, float(f[5]) if len(f) > 5 else FC[order])
            d_w = torch.empty((ns, n), dtype=torch.int32, device=dev)
            d_out = torch.zeros((ns, n // D, 2), dtype=torch.float32, device=dev)
            ctx = cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32, n_streams=ns, cuda_stream=stream.cuda_stream)
            for c in range(ns):
                ctx.synth_words(d_w[c], n, seed=c, tone_i=ti, tone_q=tq)
            ctx.set_profiling(True)
            med, mn = timeit(stream, lambda: ctx.rx_submit(d_w, n * 4, d_out, stride_bytes=n * 4, filt_stride=n // D, mem=cb.MEM_DEVICE))
            k = ctx.read_profile()
            nbytes = ns * n * (4 + 8 / D)
            fi = ctx.filter_info()
            print(json.dumps({"case": case, "ms_median": med, "ms_min": mn, "kernel_ms_avg": float(np.mean(k[-20:])),
                              "GBps": nbytes / (float(np.mean(k[-20:])) / 1e3) / 1e9, "frac": nbytes / (float(np.mean(k[-20:])) / 1e3) / 1e9 / PEAK,
                              "Gsps": ns * n / (float(np.mean(k[-20:])) / 1e3) / 1e9, "form": fi.form, "run_len": fi.run_len,
                              "window": fi.lookback_window}), flush=True)
            ctx.close(); del d_w, d_out
        elif f[0] == "cf32":
            lg, order, D = int(f[1]), int(f[2]), int(f[3])
            n = 1 << lg
            sos = cb.butter_lowpass(order, 4e6, FC[order])
            x = (torch.rand((n, 2), device=dev, dtype=torch.float32) - 0.5)
            out = torch.zeros((n // D, 2), dtype=torch.float32, device=dev)
            ctx = cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32, cuda_stream=stream.cuda_stream)
            ctx.set_profiling(True)
            med, mn = timeit(stream, lambda: ctx.filter_submit(cb.IN_CF32, x, n, out))
            k = ctx.read_profile()
            nbytes = n * (8 + 8 / D)
            ka = float(np.mean(k[-20:]))
            print(json.dumps({"case": case, "ms_median": med, "ms_min": mn, "kernel_ms_avg": ka, "GBps": nbytes / (ka / 1e3) / 1e9,
                              "frac": nbytes / (ka / 1e3) / 1e9 / PEAK, "Gsps": n / (ka / 1e3) / 1e9}), flush=True)
            ctx.close(); del x, out
        elif f[0] == "unpack":
            lg, B = int(f[1]), int(f[2])
            n = 1 << lg
            d_w = torch.empty(n, dtype=torch.int32, device=dev)
            ctx = cb.Context(interleave_block_words=B, cuda_stream=stream.cuda_stream)
            ctx.synth_words(d_w, n, tone_i=ti, tone_q=tq)
            m = n if B == 0 else n // 2
            bufs = dict(cs16=torch.zeros((m, 2), dtype=torch.int16, device=dev), meta=torch.zeros(m, dtype=torch.uint8, device=dev),
                        cf32=torch.zeros((m, 2), dtype=torch.float32, device=dev))
            if B:
                bufs.update(cs16_b=torch.zeros((m, 2), dtype=torch.int16, device=dev), meta_b=torch.zeros(m, dtype=torch.uint8, device=dev),
                            cf32_b=torch.zeros((m, 2), dtype=torch.float32, device=dev))
            ctx.set_profiling(True)
            med, mn = timeit(stream, lambda: ctx.unpack_submit(d_w, **bufs))
            k = ctx.read_profile()
            ka = float(np.mean(k[-20:]))
            nbytes = n * 17
            print(json.dumps({"case": case, "ms_median": med, "ms_min": mn, "kernel_ms_avg": ka, "GBps": nbytes / (ka / 1e3) / 1e9,
                              "frac": nbytes / (ka / 1e3) / 1e9 / PEAK, "Gsps": n / (ka / 1e3) / 1e9}), flush=True)
            ctx.close(); del d_w, bufs
        elif f[0] == "tx":
            lg, fmt = int(f[1]), f[2]
            n = 1 << lg
            dt, es = {"CS16": (torch.int16, 4), "CS8": (torch.int8, 2), "CF32": (torch.float32, 8), "CF64": (torch.float64, 16)}[fmt]
            x = (torch.rand((n, 2), device=dev) * 2 - 1)
            x = (x * (4095 if fmt == "CS16" else 127)).to(dt) if fmt in ("CS16", "CS8") else x.to(dt)
            out = torch.zeros(n, dtype=torch.int32, device=dev)
            ctx = cb.Context(cuda_stream=stream.cuda_stream)
            ctx.set_profiling(True)
            med, mn = timeit(stream, lambda: ctx.tx_pack_submit(getattr(cb, "FMT_" + fmt), x, n, out))
            k = ctx.read_profile()
            ka = float(np.mean(k[-20:]))
            nbytes = n * (es + 4)
            print(json.dumps({"case": case, "ms_median": med, "ms_min": mn, "kernel_ms_avg": ka, "GBps": nbytes / (ka / 1e3) / 1e9,
                              "frac": nbytes / (ka / 1e3) / 1e9 / PEAK, "Gsps": n / (ka / 1e3) / 1e9}), flush=True)
            ctx.close(); del x, out


if __name__ == "__main__":
    main()
