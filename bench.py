#!/usr/bin/env python
"""bench.py - throughput of the RX sample path (raw SMI words -> filtered, decimated CF32).

Workload (BASELINE.json configs[3], "C4"): per GPU a batch of 64 independent 2^24-word captures,
fused unpack -> order-4 Butterworth low-pass (fc = 250 kHz @ 4 MHz, iir1 design) -> decimate by 8,
CF32 out.  Captures are independent, so ranks share nothing (weak scaling, no collective on the
data path).  One "step" = one pass of the path over the rank's batch.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Prints ONE JSON line on rank 0.  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "Msamples/s raw SMI->filtered CF32"
UNIT = "Msamples/s"
FILTER = (4, 4e6, 250e3)     # order, fs, fc  (the config-3/4 filter)
DECIM = 8
BYTES_PER_SAMPLE = 4 + 8 / DECIM     # algorithmic: one SMI word in, 8/D bytes of CF32 out
FALLBACK_HBM_GBS = 6650.0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--captures", type=int, default=64, help="independent captures per GPU")
    ap.add_argument("--log2-words", type=int, default=24, help="log2 of the words per capture")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the host-buffer leg (0: min(steps, 5))")
    ap.add_argument("--cpu-log2-words", type=int, default=22, help="log2 words per capture in the CPU legs")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg")
    return ap.parse_args()


def workload_name(a):
    return ("C4: %d independent 2^%d-word captures per GPU, fused unpack->Butterworth order-%d LPF "
            "(fc=%g kHz @ %g MHz)->decimate-by-%d, CF32 out" %
            (a.captures, a.log2_words, FILTER[0], FILTER[2] / 1e3, FILTER[1] / 1e6, DECIM))


# ------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = str(gpu_index)
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8 or f[0] != self.idx:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smax), "power_w_max": max(power),
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------ helpers
def env_rank():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def measured_peak():
    """HBM copy ceiling of this pool's B200s (driver-written MEASURED_PEAKS.json), else the fallback the
    profiling recipe states."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except (OSError, KeyError, ValueError):
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic_per_launch(a):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel from the committed ncu --set full
    capture (profiles/roofline_traffic.json, written by profiles/summarise.py), if it matches this workload."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            t = json.load(f)
        if t.get("captures") == a.captures and t.get("log2_words") == a.log2_words:
            return float(t["dram_bytes_per_launch"])
    except (OSError, KeyError, ValueError):
        pass
    return None


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


# ------------------------------------------------------------------ CPU legs (reference build / oracle port)
class CpuPath:
    """The reference's own CPU implementation of the path on `threads` host threads, one independent
    capture per thread (how C4 parallelises on a CPU; a single stream's IIR is serial in the reference).
    kind = "reference": oracle/_ref (the unmodified caribou_smi_read + iir1 compiled from /root/reference);
    kind = "port": the oracle restatement, used only when the reference build is absent."""

    def __init__(self, log2_words, threads):
        from oracle import pyoracle as po
        self.po = po
        self.n = 1 << log2_words
        self.threads = threads
        self.kind = "reference" if po.ref_available() else "port"
        self.sos = po.butter_lowpass_sos(*FILTER)
        distinct = min(threads, 8)
        caps = [po.gen_words(self.n, seed=s) for s in range(distinct)]
        self.caps = [caps[t % distinct] for t in range(threads)]
        if self.kind == "reference":
            self.paths = [po.RefRxPath(c, FILTER[0], FILTER[1], FILTER[2], DECIM) for c in self.caps]

    def _one(self, t):
        if self.kind == "reference":
            p = self.paths[t]
            p.rewind()
            return p.run().shape[0]
        return self.po.rx_path_f32(self.caps[t], self.sos, DECIM).shape[0]

    def step(self):
        """One pass: every thread filters its capture once.  Returns (seconds, samples)."""
        out = [0] * self.threads
        def work(t):
            out[t] = self._one(t)
        ths = [threading.Thread(target=work, args=(t,)) for t in range(self.threads)]
        t0 = time.perf_counter()
        for th in ths:
            th.start()
        for th in ths:
            th.join()
        dt = time.perf_counter() - t0
        assert all(m == self.n // DECIM for m in out), out
        return dt, self.threads * self.n

    def sample_text(self):
        return ("%d threads x one 2^%d-word capture each per pass (same generator, filter and decimation as "
                "the GPU workload; unpack via caribou_smi_read on a memfd, iir1 DF-II cascades)"
                % (self.threads, int(np.log2(self.n))))


def cpu_baseline_leg(a, budget_s=12.0):
    threads = host_threads()
    cp = CpuPath(a.cpu_log2_words, threads)
    cp.step()                                   # warm-up (page faults, caches)
    tot_t = tot_n = 0.0
    reps = 0
    while tot_t < budget_s and reps < 64:
        dt, n = cp.step()
        tot_t += dt; tot_n += n; reps += 1
    return {"value": tot_n / tot_t / 1e6, "unit": UNIT, "cores": threads, "kind": cp.kind,
            "sample": "%d passes of %s" % (reps, cp.sample_text())}


def run_reference(a):
    rank, _, world = env_rank()
    if rank != 0:
        return 0
    threads = host_threads()
    cp = CpuPath(a.cpu_log2_words, threads)
    for _ in range(max(1, a.warmup)):
        cp.step()
    tot_t = tot_n = 0.0
    for _ in range(a.steps):
        dt, n = cp.step()
        tot_t += dt; tot_n += n
    v = tot_n / tot_t / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": tot_t / a.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a), "step": "bounded sample: " + cp.sample_text()},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": cp.kind,
                             "sample": "%d passes of %s" % (a.steps, cp.sample_text())},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------ the B200 arm
def run_b200(a):
    import torch
    import torch.distributed as dist
    import cariboulite_b200 as cb
    from oracle import pyoracle as po      # only for the tone tables / filter design check and the cpu_baseline leg

    rank, local_rank, world = env_rank()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: libcaribou_cuda has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    cb.lib()

    ns, n = a.captures, 1 << a.log2_words
    nout = n // DECIM
    sos = cb.butter_lowpass(*FILTER)
    ti, tq = po.tone_tables()
    # a stream of our own, made torch's current one: the library launches on it and torch.cuda.Event
    # (which only sees torch's current stream) times it
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)

    d_w = torch.empty((ns, n), dtype=torch.int32, device=dev)
    d_out = torch.zeros((ns, nout, 2), dtype=torch.float32, device=dev)
    ctx = cb.Context(device=local_rank, sos=sos, decimation=DECIM, parity=cb.PARITY_F32, out_format=cb.FMT_CF32,
                     n_streams=ns, cuda_stream=stream.cuda_stream)
    for c in range(ns):      # capture c of rank r: seed r * ns + c, device-generated in place
        ctx.synth_words(d_w[c], n, seed=rank * ns + c, tone_i=ti, tone_q=tq)
    torch.cuda.synchronize(dev)

    def step():
        ctx.rx_submit(d_w, n * 4, d_out, stride_bytes=n * 4, filt_stride=nout, mem=cb.MEM_DEVICE)

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident leg: `value` and the roofline
    for _ in range(max(3, a.warmup)):
        step()
    ret, res = ctx.wait()
    assert ret == n and res.n_filtered == nout and res.n_misaligned_chunks == 0, (ret, res.n_filtered)
    ctx.set_profiling(True)
    sampler = ClockSampler(local_rank if "CUDA_VISIBLE_DEVICES" not in os.environ else
                           os.environ["CUDA_VISIBLE_DEVICES"].split(",")[local_rank])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = ctx.launch_count
    barrier(); torch.cuda.synchronize(dev)
    if rank == 0:
        sampler.start()
    e0.record(stream)
    for _ in range(a.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize(dev); barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = ctx.launch_count - l0
    kern_ms = ctx.read_profile()
    ctx.set_profiling(False)
    # keep the GPU busy a little longer if the timed region was too short for nvidia-smi to sample it
    clocks = None
    if rank == 0:
        t_end = time.time() + max(0.0, 1.2 - ms_total / 1e3)
        while time.time() < t_end:
            step()
        ctx.wait()
        clocks = sampler.stop()
    ret, res = ctx.wait()
    assert ret == n
    barrier()

    samples_per_step = ns * n * world
    value = samples_per_step * a.steps / (ms_total / 1e3) / 1e6
    peak, peak_src = measured_peak()
    k_avg = float(np.mean(kern_ms)) if len(kern_ms) else float("nan")
    alg_bytes = ns * n * BYTES_PER_SAMPLE
    achieved = alg_bytes / (k_avg / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": "iir_scan_kernel<K=2,D=8,IN_WORDS,PAR,Y32> (fused unpack+IIR+decimate, float-staged)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic_per_launch(a), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "bytes_per_sample": BYTES_PER_SAMPLE,
                "kernel_ms_avg": k_avg, "kernel_ms_min": float(np.min(kern_ms)) if len(kern_ms) else None,
                "kernel_launches_timed": int(len(kern_ms)),
                "frac_of_spec_8000": achieved / 8000.0}

    # ---- host-buffer leg through the C ABI: H2D of the raw words and D2H of the filtered output inside
    e2e = None
    if not a.no_e2e:
        k2 = a.e2e_steps or min(a.steps, 5)
        h_w = torch.empty((ns, n), dtype=torch.int32).pin_memory()
        h_out = torch.zeros((ns, nout, 2), dtype=torch.float32).pin_memory()
        h_w.copy_(d_w)
        torch.cuda.synchronize(dev)
        def step_host():
            ctx.rx_submit(h_w, n * 4, h_out, stride_bytes=n * 4, filt_stride=nout, mem=cb.MEM_HOST_PINNED)
        ctx.reset()
        for _ in range(2):
            step_host()
        ctx.wait()
        barrier(); torch.cuda.synchronize(dev)
        e0.record(stream)
        for _ in range(k2):
            step_host()
        e1.record(stream)
        torch.cuda.synchronize(dev); barrier()
        ret, res = ctx.wait()
        assert ret == n
        ms2 = max_over_ranks(e0.elapsed_time(e1))
        # the result read back is the filtered stream itself; check it is the one the device leg produces
        ctx.reset(); step_host(); ctx.wait()
        ctx.reset(); step(); ctx.wait()
        assert torch.equal(h_out.to(dev), d_out), "host-buffer leg and device leg disagree"
        e2e = {"value": samples_per_step * k2 / (ms2 / 1e3) / 1e6, "unit": UNIT, "steps": k2,
               "h2d_bytes_per_step": ns * n * 4, "d2h_bytes_per_step": ns * nout * 8,
               "ms_per_step": ms2 / k2, "api": "caribou_cuda_rx_submit(mem=HOST_PINNED) + caribou_cuda_wait"}
        del h_w, h_out

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu:
        cpu = cpu_baseline_leg(a)

    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(3, a.warmup),
            "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a), "captures_per_gpu": ns, "words_per_capture": n,
                       "parity_mode": "F32", "chunk_bytes": 524288,
                       "l2": "inputs larger than L2 (%.1f GiB read per step per GPU, no reuse between steps)" % (ns * n * 4 / 2**30),
                       "sharding": "independent captures per rank, no data-path collective"},
            "roofline": roofline, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)
    return 0


def main():
    a = parse_args()
    _, _, world = env_rank()
    if world != a.gpus and world > 1:
        a.gpus = world
    if a.gpus > 1 and world == 1 and a.impl == "b200":
        # convenience: relaunch under torchrun when called directly with --gpus N
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(a.gpus),
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29531")] + sys.argv
        return subprocess.call(cmd)
    return run_reference(a) if a.impl == "reference" else run_b200(a)


if __name__ == "__main__":
    sys.exit(main())
