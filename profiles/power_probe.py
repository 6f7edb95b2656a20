#!/usr/bin/env python
"""Is the fused kernel's time set by the SM clock the power cap leaves?  Times the C4 launch (a) back to back,
(b) one launch every 30 ms, (c) the first launches after 2 s of idling, with nvidia-smi sampled every 50 ms, and
does the same for a plain device copy.  One JSON line per regime.

    python profiles/power_probe.py [order] [D]
"""
import json, os, subprocess, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import cariboulite_b200 as cb
from profiles.probe import tone_tables, FC


class Smi:
    def __init__(self):
        self.lines = []
        self.p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap",
                                   "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE, text=True)
        threading.Thread(target=self._pump, daemon=True).start()

    def _pump(self):
        for ln in self.p.stdout:
            self.lines.append((time.time(), ln))

    def window(self, t0, t1):
        sm, pw, cap = [], [], 0
        for t, ln in self.lines:
            if t0 <= t <= t1:
                f = [x.strip() for x in ln.split(",")]
                try:
                    sm.append(float(f[0])); pw.append(float(f[1])); cap += f[2].lower().startswith("active")
                except (ValueError, IndexError):
                    pass
        if not sm:
            return {}
        return {"sm_mhz_median": float(np.median(sm)), "sm_mhz_min": min(sm), "power_w_mean": float(np.mean(pw)),
                "power_w_max": max(pw), "power_cap_samples": cap, "samples": len(sm)}


def regimes(name, smi, launch, read_times, sync):
    # (a) back to back
    for _ in range(5):
        launch()
    sync(); read_times()
    t0 = time.time()
    batches = []
    for _ in range(4):                      # the timing ring holds 256 launches: read it every 200
        for _ in range(200):
            launch()
        batches.append(read_times())
    t1 = time.time()
    print(json.dumps({"what": name, "regime": "back to back 4 x 200", "kernel_ms_first10": float(np.mean(batches[0][:10])),
                      "kernel_ms_batch_means": [round(float(np.mean(b)), 4) for b in batches],
                      **smi.window(t0 + 0.2, t1)}), flush=True)
    # (b) one launch every 30 ms
    time.sleep(2.0)
    t0 = time.time()
    for _ in range(60):
        launch(); sync(); time.sleep(0.03)
    t1 = time.time()
    k = read_times()
    print(json.dumps({"what": name, "regime": "one launch per 30 ms x60", "kernel_ms_mean": float(np.mean(k[-50:])),
                      "kernel_ms_min": float(np.min(k[-50:])), **smi.window(t0, t1)}), flush=True)
    # (c) bursts of 8 after idling
    res = []
    for _ in range(3):
        time.sleep(2.0)
        for _ in range(8):
            launch()
        sync()
        res.append([round(float(x), 4) for x in read_times()[-8:]])
    print(json.dumps({"what": name, "regime": "8 launches after 2 s idle, x3", "kernel_ms": res}), flush=True)


def main():
    order = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    D = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    smi = Smi()
    ns, n = 64, 1 << 24
    ti, tq = tone_tables()
    sos = cb.butter_lowpass(order, 4e6, FC[order])
    d_w = torch.empty((ns, n), dtype=torch.int32, device=dev)
    d_out = torch.zeros((ns, n // D, 2), dtype=torch.float32, device=dev)
    ctx = cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32, n_streams=ns, cuda_stream=stream.cuda_stream)
    for c in range(ns):
        ctx.synth_words(d_w[c], n, seed=c, tone_i=ti, tone_q=tq)
    ctx.wait()
    ctx.set_profiling(True)
    regimes("fused order %d D %d (%.3f GB algorithmic)" % (order, D, ns * n * (4 + 8 / D) / 1e9), smi,
            lambda: ctx.rx_submit(d_w, n * 4, d_out, stride_bytes=n * 4, filt_stride=n // D, mem=cb.MEM_DEVICE),
            lambda: np.asarray(ctx.read_profile(), dtype=np.float64), lambda: ctx.wait())
    ctx.close()
    # plain copy of the same number of bytes moved (read 4 GiB + write 1 GiB ~ copy of 2.5 GiB)
    a = d_w.view(-1)[: (5 << 30) // 8]
    b = torch.empty_like(a)
    ev = []

    def launch():
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); b.copy_(a); e1.record(stream); ev.append((e0, e1))

    def read_times():
        torch.cuda.synchronize()
        t = [x.elapsed_time(y) for x, y in ev]
        ev.clear()
        return np.asarray(t)

    regimes("torch copy of 2.5 GiB (5.37 GB moved)", smi, launch, read_times, torch.cuda.synchronize)
    smi.p.terminate()


if __name__ == "__main__":
    main()
