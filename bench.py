#!/usr/bin/env python
"""bench.py - throughput of the RX sample path (raw SMI words -> filtered, decimated CF32).

Workload of the headline line (BASELINE.json configs[3], "C4"): per GPU a batch of 64 independent 2^24-word
captures, fused unpack -> order-4 Butterworth low-pass (fc = 250 kHz @ 4 MHz, iir1 design) -> decimate by 8,
CF32 out.  Captures are independent, so ranks share nothing (weak scaling, no collective on the data path).
One "step" = one pass of the path over the rank's batch.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Prints ONE JSON line on rank 0.  Besides the contract's keys it carries
  roofline            the fused scan kernel: burst figure over the K timed steps, and `sustained` = the same kernel
                      over >= 2 s of back-to-back steps with the clocks sampled over exactly that window
  e2e                 host buffers through caribou_cuda_rx_submit(HOST_PINNED) + wait, with the measured copy
                      ceiling of this box at this N beside it (`ceiling_msps`)
  c5                  BASELINE.json configs[4]: ONE 2^33-word stream time-chunked over the ranks with the IIR state
                      handed over on the device (caribou_cuda_shard_submit); at N = 1 eight carried calls
  configs             (N = 1) C1, C2, C3 kernel times against their algorithmic bytes, the reference's 1-thread CPU
                      rates BASELINE.md section 3 promises, and the Soapy-shaped read loop on both sides (e2e_shim)
  cpu_baseline        (N = 1) the reference's own CPU code on all host threads, bounded sample
See DESIGN.md "Measurement" for every field.
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
MTU = 131072


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--captures", type=int, default=64, help="independent captures per GPU")
    ap.add_argument("--log2-words", type=int, default=24, help="log2 of the words per capture")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the host-buffer leg (0: min(steps, 5))")
    ap.add_argument("--cpu-log2-words", type=int, default=22, help="log2 words per capture in the cpu_baseline leg")
    ap.add_argument("--sustain-seconds", type=float, default=2.0, help="length of the sustained-clock window")
    ap.add_argument("--c5-log2-words", type=int, default=33, help="log2 of the words of the time-chunked stream")
    ap.add_argument("--no-cpu", action="store_true", help="skip every CPU leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg")
    ap.add_argument("--no-c5", action="store_true", help="skip the time-chunked stream leg")
    ap.add_argument("--no-configs", action="store_true", help="skip the per-config records")
    ap.add_argument("--no-sustain", action="store_true", help="skip the sustained-clock window")
    return ap.parse_args()


def workload_name(a):
    return ("C4: %d independent 2^%d-word captures per GPU, fused unpack->Butterworth order-%d LPF "
            "(fc=%g kHz @ %g MHz)->decimate-by-%d, CF32 out" %
            (a.captures, a.log2_words, FILTER[0], FILTER[2] / 1e3, FILTER[1] / 1e6, DECIM))


def config_of(a):
    n = 1 << a.log2_words
    return {"workload": workload_name(a), "captures_per_gpu": a.captures, "words_per_capture": n,
            "parity_mode": "F32", "chunk_bytes": 524288,
            "l2": "inputs larger than L2 (%.1f GiB read per step per GPU, no reuse between steps)" % (a.captures * n * 4 / 2**30),
            "sharding": "independent captures per rank, no data-path collective"}


# ------------------------------------------------------------------ clocks
class ClockSampler:
    """SM clock, power and throttle reasons of one GPU sampled from a thread through NVML every few milliseconds
    (nvidia-smi's 200 ms polling cannot see a 20 ms timed region)."""
    NAMES = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}

    def __init__(self, gpu_index, period_s=0.004):
        self.period = period_s
        self.samples = []
        self.stop_flag = threading.Event()
        self.h = None
        self.nv = None
        self.smax = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(int(gpu_index))
            self.smax = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _read(self):
        nv = self.nv
        sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
        try:
            pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
        except Exception:
            pw = float("nan")
        rs = 0
        for fn in ("nvmlDeviceGetCurrentClocksEventReasons", "nvmlDeviceGetCurrentClocksThrottleReasons"):
            f = getattr(nv, fn, None)
            if f is None:
                continue
            try:
                rs = f(self.h)
                break
            except Exception:
                continue
        return sm, pw, rs

    def _loop(self):
        while not self.stop_flag.is_set():
            try:
                self.samples.append(self._read())
            except Exception:
                pass
            time.sleep(self.period)

    def start(self):
        self.samples = []
        self.stop_flag.clear()
        if self.nv is None:
            return
        self.t = threading.Thread(target=self._loop, daemon=True)
        self.t.start()

    def stop(self):
        if self.nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["NVML unavailable"]}
        self.stop_flag.set()
        self.t.join(timeout=2)
        if not self.samples:
            try:
                self.samples.append(self._read())
            except Exception:
                return {"sm_mhz": None, "sm_max_mhz": self.smax, "reasons": ["no samples"]}
        sm = [s[0] for s in self.samples]
        pw = [s[1] for s in self.samples if s[1] == s[1]]
        bits = 0
        for s in self.samples:
            bits |= int(s[2])
        reasons = sorted(n for n, b in self.NAMES.items() if bits & b)
        return {"sm_mhz": float(statistics.median(sm)), "sm_min_mhz": float(min(sm)), "sm_max_mhz": float(self.smax),
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons, "source": "NVML"}


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


def bind_to_gpu_numa_node(local_rank):
    """Pin this rank's threads to the CPUs local to its GPU (and, through first touch, its page-locked buffers to
    that node's memory) before anything is allocated.  Returns a description for the bench line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        idx = local_rank
        if "CUDA_VISIBLE_DEVICES" in os.environ:
            idx = int(os.environ["CUDA_VISIBLE_DEVICES"].split(",")[local_rank])
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dom, rest = bus.split(":", 1)
        path = "/sys/bus/pci/devices/%s:%s" % (dom[-4:].lower(), rest.lower())
        node = int(open(path + "/numa_node").read().strip())
        cpus = set()
        for part in open(path + "/local_cpulist").read().strip().split(","):
            if "-" in part:
                lo, hi = part.split("-")
                cpus.update(range(int(lo), int(hi) + 1))
            elif part:
                cpus.add(int(part))
        allowed = set(os.sched_getaffinity(0))
        use = sorted(cpus & allowed)
        if use and len(use) < len(allowed):
            os.sched_setaffinity(0, use)
        return {"numa_node": node, "local_cpus": len(cpus), "bound_cpus": len(use) if use else len(allowed)}
    except Exception as e:      # no sysfs entry (container), NVML missing ...: run unbound and say so
        return {"numa_node": None, "note": "unbound (%s)" % type(e).__name__}


# ------------------------------------------------------------------ CPU legs (reference build / oracle port)
class CpuPath:
    """The reference's own CPU implementation of the path on `threads` host threads over `captures` independent
    captures of 2^log2_words words (how C4 parallelises on a CPU; a single stream's IIR is serial in the
    reference).  kind = "reference": oracle/_ref (the unmodified caribou_smi_read + iir1 compiled from
    /root/reference); kind = "port": the oracle restatement, used only when the reference build is absent."""

    def __init__(self, log2_words, threads, captures=None, seed0=0):
        from oracle import pyoracle as po
        self.po = po
        self.n = 1 << log2_words
        self.threads = threads
        self.captures = captures or threads
        self.kind = "reference" if po.ref_available() else "port"
        self.sos = po.butter_lowpass_sos(*FILTER)
        if self.kind == "reference":
            self.paths = [po.RefRxPath(po.gen_words(self.n, seed=seed0 + c), FILTER[0], FILTER[1], FILTER[2], DECIM)
                          for c in range(self.captures)]
        else:
            self.caps = [po.gen_words(self.n, seed=seed0 + c) for c in range(self.captures)]

    def _one(self, c):
        if self.kind == "reference":
            p = self.paths[c]
            p.rewind()
            return p.run().shape[0]
        return self.po.rx_path_f32(self.caps[c], self.sos, DECIM).shape[0]

    def step(self):
        """One pass: every capture is filtered once, `threads` at a time.  Returns (seconds, samples)."""
        out = [0] * self.captures
        nxt = [0]
        lock = threading.Lock()

        def work():
            while True:
                with lock:
                    c = nxt[0]
                    nxt[0] += 1
                if c >= self.captures:
                    return
                out[c] = self._one(c)
        ths = [threading.Thread(target=work) for _ in range(self.threads)]
        t0 = time.perf_counter()
        for th in ths:
            th.start()
        for th in ths:
            th.join()
        dt = time.perf_counter() - t0
        assert all(m == self.n // DECIM for m in out), out
        return dt, self.captures * self.n

    def sample_text(self):
        return ("%d independent 2^%d-word captures per pass on %d threads (same generator, filter and decimation as "
                "the GPU workload; unpack via caribou_smi_read on a memfd, iir1 DF-II cascades)"
                % (self.captures, int(np.log2(self.n)), self.threads))


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
    """The reference arm: the reference's own CPU code on the SAME config as the B200 arm - every step filters the
    rank's whole batch (64 captures of 2^24 words, seeds as on the GPU) on all host threads."""
    rank, _, world = env_rank()
    if rank != 0:
        return 0
    threads = host_threads()
    cp = CpuPath(a.log2_words, threads, captures=a.captures)
    for _ in range(max(1, min(a.warmup, 2))):
        cp.step()
    tot_t = tot_n = 0.0
    for _ in range(a.steps):
        dt, n = cp.step()
        tot_t += dt; tot_n += n
    v = tot_n / tot_t / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": tot_t / a.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config_of(a),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": cp.kind,
                             "sample": "%d passes of %s" % (a.steps, cp.sample_text())},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def cpu_single_thread_legs(po):
    """BASELINE.md section 3 runs 1 and 2 on one host thread, the reference's own object code (oracle/_ref):
    caribou_smi_read over a memfd for C1 with and without the read(2) copy, and Stream::ReadSamplesGen(CF32) with
    the three filters the Soapy layer ships (order 6 at fc = 10 / 25 / 50 kHz) on a 2^24-sample slice."""
    if not po.ref_available():
        return {"unavailable": "oracle/_ref is missing"}
    out = {}
    n = 1 << 22
    dev = po.RefDev(po.gen_words(n))
    cs16 = np.zeros((n + 1, 2), dtype=np.int16)
    meta = np.zeros(n + 1, dtype=np.uint8)
    t_read, t_copy = [], []
    for rep in range(6):
        dev.rewind()
        t0 = time.perf_counter()
        r = dev.smi_read_into(cs16, meta, n)
        t1 = time.perf_counter()
        assert r == n
        dev.rewind()
        t2 = time.perf_counter()
        assert dev.read_only(n) == n
        t3 = time.perf_counter()
        if rep:
            t_read.append(t1 - t0); t_copy.append(t3 - t2)
    dev.close()
    tr, tc = min(t_read), min(t_copy)
    out["c1_smi_read"] = {"what": "caribou_smi_read (caribou_smi.c:632-682) -> CS16 + meta, 2^22 words from a memfd, best of 5",
                          "msps_with_read_copy": n / tr / 1e6, "ms": tr * 1e3, "read_copy_ms": tc * 1e3,
                          "msps_without_read_copy": n / max(tr - tc, 1e-9) / 1e6}
    n = 1 << 24
    w = po.gen_words(n)
    buf = np.zeros((MTU + 1, 2), dtype=np.float32)
    for t, name in ((1, "fc10k"), (2, "fc25k"), (3, "fc50k")):
        dev = po.RefDev(w)
        dev.stream_open()
        dev.stream_set_format("CF32")
        dev.stream_set_filter(t)
        got = 0
        t0 = time.perf_counter()
        while got < n:
            r = dev.stream_read_into(buf, MTU)
            if r <= 0:
                break
            got += r
        dt = time.perf_counter() - t0
        dev.close()
        assert got == n
        out["stream_cf32_order6_" + name] = {"msps": n / dt / 1e6, "ms_per_mtu": dt / (n / MTU) * 1e3}
    out["what_stream"] = "Stream::ReadSamplesGen(CF32) (CaribouliteStream.cpp:304-323,370-382), 2^24 words in MTU reads, 1 thread"
    return out


# ------------------------------------------------------------------ the B200 arm
def run_b200(a):
    import torch
    import torch.distributed as dist

    rank, local_rank, world = env_rank()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: libcaribou_cuda has no CPU fallback")
    numa = bind_to_gpu_numa_node(local_rank)        # before any allocation: page-locked buffers land on the local node
    import cariboulite_b200 as cb
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    cb.lib()
    gpu_index = local_rank
    try:
        if os.environ.get("CUDA_VISIBLE_DEVICES"):
            gpu_index = int(os.environ["CUDA_VISIBLE_DEVICES"].split(",")[local_rank])
    except (ValueError, IndexError):
        gpu_index = local_rank

    ns, n = a.captures, 1 << a.log2_words
    nout = n // DECIM
    sos = cb.butter_lowpass(*FILTER)
    ti, tq = cb.tone_tables()
    # a stream of our own, made torch's current one: the library launches on it and torch.cuda.Event
    # (which only sees torch's current stream) times it
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    peak, peak_src = measured_peak()

    d_w = torch.empty((ns, n), dtype=torch.int32, device=dev)
    d_out = torch.zeros((ns, nout, 2), dtype=torch.float32, device=dev)
    ctx = cb.Context(device=local_rank, sos=sos, decimation=DECIM, parity=cb.PARITY_F32, out_format=cb.FMT_CF32,
                     n_streams=ns, cuda_stream=stream.cuda_stream)
    for c in range(ns):      # capture c of rank r: seed r * ns + c, device-generated in place
        ctx.synth_words(d_w[c], n, seed=rank * ns + c, tone_i=ti, tone_q=tq)
    torch.cuda.synchronize(dev)

    def step():
        ctx.rx_submit(d_w, n * 4, d_out, stride_bytes=n * 4, filt_stride=nout, mem=cb.MEM_DEVICE)

    # ---- device-resident leg: `value` and the burst roofline
    for _ in range(max(3, a.warmup)):
        step()
    ret, res = ctx.wait()
    assert ret == n and res.n_filtered == nout and res.n_misaligned_chunks == 0, (ret, res.n_filtered)
    ctx.set_profiling(True)
    sampler = ClockSampler(gpu_index)
    l0 = ctx.launch_count
    barrier(); torch.cuda.synchronize(dev)
    if rank == 0:
        sampler.start()
    e0.record(stream)
    for _ in range(a.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize(dev); barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = ctx.launch_count - l0
    kern_ms = ctx.read_profile()
    ret, res = ctx.wait()
    assert ret == n

    samples_per_step = ns * n * world
    value = samples_per_step * a.steps / (ms_total / 1e3) / 1e6
    k_avg = float(np.mean(kern_ms)) if len(kern_ms) else float("nan")
    alg_bytes = ns * n * BYTES_PER_SAMPLE
    achieved = alg_bytes / (k_avg / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": "iir_scan_kernel<K=2,D=8,IN_WORDS,PAR,Y32> (fused unpack+IIR+decimate, float-staged)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic_per_launch(a), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "bytes_per_sample": BYTES_PER_SAMPLE,
                "kernel_ms_avg": k_avg, "kernel_ms_min": float(np.min(kern_ms)) if len(kern_ms) else None,
                "kernel_launches_timed": int(len(kern_ms)),
                "frac_of_spec_8000": achieved / 8000.0,
                "regime": "burst: the %d timed steps (%.0f ms) start on a GPU that idled through the set-up" % (a.steps, ms_total)}

    # ---- sustained window: the same step back to back until the power management has settled; the clock sampler
    # covers exactly this window and the kernel time is the mean of the last launches inside it
    if not a.no_sustain:
        n_sus = max(64, int(a.sustain_seconds / max(ms_total / a.steps / 1e3, 1e-5)))
        barrier(); torch.cuda.synchronize(dev)
        ctx.read_profile()
        s2 = ClockSampler(gpu_index, period_s=0.02)
        if rank == 0:
            s2.start()
        e0.record(stream)
        for i in range(n_sus):
            step()
            if (i & 31) == 31:
                ctx.wait()            # keeps the queue of optimistic submits short; costs one launch gap per 32 steps
        e1.record(stream)
        torch.cuda.synchronize(dev)
        sus_ms = e0.elapsed_time(e1)
        c2 = s2.stop() if rank == 0 else None
        km = ctx.read_profile()                 # the last <= 256 launches of the window
        ks = float(np.mean(km)) if len(km) else float("nan")
        ach_s = alg_bytes / (ks / 1e3) / 1e9
        roofline["sustained"] = {"seconds": sus_ms / 1e3, "steps": n_sus, "kernel_ms_avg": ks,
                                 "kernel_launches_averaged": int(len(km)), "achieved": ach_s, "frac": ach_s / peak,
                                 "ms_per_step": sus_ms / n_sus, "clocks": c2}
        barrier()
    ctx.set_profiling(False)

    # ---- host-buffer leg through the C ABI: H2D of the raw words and D2H of the filtered output inside
    e2e = None
    if not a.no_e2e:
        k2 = a.e2e_steps or min(a.steps, 5)
        h_w = torch.empty((ns, n), dtype=torch.int32).pin_memory()
        h_out = torch.zeros((ns, nout, 2), dtype=torch.float32).pin_memory()
        h_w.copy_(d_w)
        torch.cuda.synchronize(dev)

        # what this box lets N ranks copy at once: the same bytes as a step moves, H2D and D2H side by side on two
        # streams, nothing else - the ceiling the host-buffer step is held against
        s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        ce0, ce1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        cdn = torch.cuda.Event(enable_timing=True)

        def bare_copies(reps):
            barrier(); torch.cuda.synchronize(dev)
            ce0.record(s_up)
            s_dn.wait_event(ce0)
            for _ in range(reps):
                with torch.cuda.stream(s_up):
                    d_w.copy_(h_w, non_blocking=True)
                with torch.cuda.stream(s_dn):
                    h_out.copy_(d_out, non_blocking=True)
            cdn.record(s_dn)
            s_up.wait_event(cdn)
            ce1.record(s_up)
            torch.cuda.synchronize(dev); barrier()
            return max_over_ranks(ce0.elapsed_time(ce1)) / reps
        bare_copies(1)
        ceil_ms = bare_copies(3)
        torch.cuda.set_stream(stream)

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
               "ms_per_step": ms2 / k2, "api": "caribou_cuda_rx_submit(mem=HOST_PINNED) + caribou_cuda_wait",
               "ceiling_ms_per_step": ceil_ms, "ceiling_msps": samples_per_step / (ceil_ms / 1e3) / 1e6,
               "frac_of_ceiling": ceil_ms / (ms2 / k2),
               "ceiling": "the same H2D + D2H bytes as bare pinned copies on two streams, all %d ranks at once" % world,
               "host_binding": numa}
        del h_w, h_out

    ctx.close()
    del d_w, d_out
    torch.cuda.empty_cache()

    # ---- C5: one long stream, time-chunked over the ranks, IIR state handed over on the device
    c5 = None
    if not a.no_c5:
        c5 = c5_leg(a, cb, torch, dist, dev, stream, rank, local_rank, world, barrier, max_over_ranks, value)

    configs = None
    cpu = None
    if rank == 0 and world == 1:
        if not a.no_configs:
            configs = config_legs(a, cb, torch, dev, stream, peak, with_cpu=not a.no_cpu)
        if not a.no_cpu:
            cpu = cpu_baseline_leg(a)

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(3, a.warmup),
            "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": config_of(a),
            "roofline": roofline, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
    if c5 is not None:
        line["c5"] = c5
    if configs is not None:
        line["configs"] = configs
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)
    return 0


def c5_leg(a, cb, torch, dist, dev, stream, rank, local_rank, world, barrier, max_over_ranks, c4_value):
    """BASELINE.json configs[4]: ONE continuous stream of 2^33 words, time-chunked over the ranks (rank g owns the
    contiguous range [g W/G, (g+1) W/G), device-generated in place), the fused path with the C4 filter (so that the
    figure compares with the C4 kernel), IIR state handed from GPU to GPU on the device (caribou_cuda_shard_submit:
    ncclAllGather on the context's stream, composition kernel, fix-up from device memory).  At N = 1 the stream goes
    through as 8 carried calls.  Parity against the sequential CPU oracle on a 2^28-sample prefix (rank 0) and on a
    window straddling every shard boundary (each rank checks the one it receives)."""
    total = 1 << a.c5_log2_words
    calls = 8 if world == 1 else 1
    n = total // (world * calls)                 # words per rank and call
    nout = n // DECIM
    sos = cb.butter_lowpass(*FILTER)
    ti, tq = cb.tone_tables()
    ctx = cb.Context(device=local_rank, sos=sos, decimation=DECIM, parity=cb.PARITY_F32, out_format=cb.FMT_CF32,
                     cuda_stream=stream.cuda_stream)
    if world > 1:
        ids = [cb.shard_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.shard_init(rank, world, ids[0])
    else:
        ctx.shard_init(0, 1)      # one rank: no communicator (the NCCL route at world = 1 is tests/test_gpu_shard.py's)
    d_w = torch.empty((calls, n), dtype=torch.int32, device=dev)
    d_out = torch.zeros((calls, nout, 2), dtype=torch.float32, device=dev)
    seed = 0xC0FFEE
    piece = 1 << 28
    for k in range(calls):
        n0 = (k * world + rank) * n              # this rank's range of call k within the one stream
        for o in range(0, n, piece):
            ctx.synth_words(d_w[k, o:o + min(piece, n - o)], min(piece, n - o), n0=n0 + o, seed=seed, tone_i=ti, tone_q=tq)
    torch.cuda.synchronize(dev)

    def one_pass():
        for k in range(calls):
            ctx.shard_submit(cb.IN_WORDS, d_w[k], n, d_out[k])

    one_pass()                                   # first pass from zero state: the one parity looks at
    assert ctx.wait()[0] == n
    lead, win = 1 << 16, 1 << 14
    npre = min(1 << 28, n)
    first_heads = [d_out[k, :win // DECIM].cpu().numpy() for k in range(calls)]
    first_prefix = d_out[0, :npre // DECIM].cpu().numpy() if rank == 0 else None
    ctx.set_profiling(True)
    one_pass()
    assert ctx.wait()[0] == n
    reps = 3
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier(); torch.cuda.synchronize(dev)
    e0.record(stream)
    for _ in range(reps):
        one_pass()
    e1.record(stream)
    torch.cuda.synchronize(dev); barrier()
    assert ctx.wait()[0] == n
    ms = max_over_ranks(e0.elapsed_time(e1)) / reps
    pass_ms, hand_own = ctx.shard_profile()
    hand_ms = max_over_ranks(hand_own)
    hand_min = -max_over_ranks(-hand_own)      # the rank that finished pass 1 last waits for nobody: the pure cost
    fix = int(ctx.fixup_samples(n))
    mode = ctx.shard_mode
    ctx.close()
    del d_w, d_out
    torch.cuda.empty_cache()

    # ---- parity (the oracle as checker): prefix on rank 0, the incoming boundary on every rank / call
    from oracle import pyoracle as po
    osos = po.butter_lowpass_sos(*FILTER)

    def rel(gv, rv):
        gv = np.asarray(gv, np.float64); rv = np.asarray(rv, np.float64)
        return float(np.max(np.sqrt(np.mean((gv - rv) ** 2, axis=0)) / np.sqrt(np.mean(rv ** 2, axis=0))))
    worst_win = 0.0
    # the oracle starts 2^16 samples before the boundary: its own start-up has died out by then
    for k in range(calls):
        g0 = (k * world + rank) * n
        if g0 == 0:
            continue
        w = po.gen_words(lead + win, n0=g0 - lead, seed=seed)
        ref = po.rx_path_f32(w, osos, DECIM)[lead // DECIM:]
        worst_win = max(worst_win, rel(first_heads[k], ref))
    if world > 1:
        t = torch.tensor([worst_win], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        worst_win = float(t.item())
    prefix = None
    if rank == 0:
        w = po.gen_words(npre, n0=0, seed=seed)
        prefix = rel(first_prefix, po.rx_path_f32(w, osos, DECIM))
    v = total / (ms / 1e3) / 1e6
    return {"workload": "C5: one %d-word stream time-chunked over %d GPU(s), %d call(s) of %d words per rank, fused "
                        "unpack->order-%d LPF->decimate-by-%d, state hand-off on the device" % (total, world, calls, n, FILTER[0], DECIM),
            "words_total": total, "words_per_rank_per_call": n, "calls": calls, "value": v, "unit": UNIT,
            "ms_per_pass": ms, "ms_per_call": ms / calls, "pass1_ms": pass_ms, "handoff_ms": hand_ms, "handoff_ms_min_over_ranks": hand_min,
            "handoff_what": "exchange of the end vectors + composition kernel + fix-up of the leading tiles: device time from the end of a "
                            "rank's pass 1; the max over ranks includes waiting for the slowest rank's pass 1, the min is the rank that waits for nobody",
            "fixup_samples": fix, "rel_rms": prefix, "rel_rms_what": "first min(2^28, shard) samples against the sequential CPU oracle",
            "worst_boundary_window": worst_win, "efficiency_vs_c4_kernel": v / c4_value,
            "shard_gib": n * 4 / 2**30, "api": "caribou_cuda_shard_submit",
            "handoff_mode": {0: "ncclAllGather", 1: "ncclSend/ncclRecv chain", 2: "peer stores over NVLink (CUDA IPC) + flags"}[mode]}


def time_launches(torch, stream, fn, sync, reps=10, warm=3):
    """Mean and best device time of `fn` (one submit) over `reps` launches, CUDA events on the launching stream."""
    for _ in range(warm):
        fn()
    sync()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for s, e in evs:
        s.record(stream)
        fn()
        e.record(stream)
    sync()
    t = [s.elapsed_time(e) for s, e in evs]
    return float(np.mean(t)), float(np.min(t))


def config_legs(a, cb, torch, dev, stream, peak, with_cpu):
    """Per-config records (N = 1): C1, C2 (B = 1 and B = 131072), C3 on the device with their algorithmic bytes
    (SURVEY 8d), the 1-thread CPU rates BASELINE.md section 3 promises, and the Soapy-shaped read loop."""
    out = {}
    ti, tq = cb.tone_tables()

    def rec(name, what, nsamp, bps, mean_ms, best_ms, extra=None):
        ach = nsamp * bps / (mean_ms / 1e3) / 1e9
        r = {"what": what, "samples": nsamp, "bytes_per_sample": bps, "ms": mean_ms, "ms_best": best_ms,
             "msps": nsamp / (mean_ms / 1e3) / 1e6, "achieved_gbs": ach, "frac": ach / peak}
        if extra:
            r.update(extra)
        out[name] = r

    # ---- C1 / C2: unpack -> CS16 + meta + CF32; one submit = alignment search + unpack + flag scan
    for name, log2n, B in (("C1", 22, 0), ("C2_B1", 26, 1), ("C2_B131072", 26, 131072)):
        n = 1 << log2n
        ctx = cb.Context(device=dev.index, interleave_block_words=B, cuda_stream=stream.cuda_stream)
        d_w = torch.empty(n, dtype=torch.int32, device=dev)
        ctx.synth_words(d_w, n, tone_i=ti, tone_q=tq)
        m = n if B == 0 else n // 2
        bufs = {}
        for suf in (("",) if B == 0 else ("", "_b")):
            bufs["cs16" + suf] = torch.zeros((m, 2), dtype=torch.int16, device=dev)
            bufs["meta" + suf] = torch.zeros(m, dtype=torch.uint8, device=dev)
            bufs["cf32" + suf] = torch.zeros((m, 2), dtype=torch.float32, device=dev)
        ctx.set_profiling(True)
        mean_ms, best_ms = time_launches(torch, stream, lambda: ctx.unpack_submit(d_w, **bufs), lambda: ctx.wait())
        km = ctx.read_profile()
        what = ("2^%d words, %s -> CS16 + meta + CF32, whole submit (alignment search, unpack, flag scan)"
                % (log2n, "S1G only" if B == 0 else "S1G/HiF interleaved B = %d, demux" % B))
        rec(name, what, n, 17.0, mean_ms, best_ms,
            {"unpack_kernels_ms": float(np.mean(km)) if len(km) else None,
             "note": "launch-latency-bound at this size" if log2n <= 22 else None})
        ctx.close()
        del d_w, bufs
    # ---- C3: 2^30 CF32 samples -> order-4 LPF -> decimate by 8 (standalone kernel (b))
    n = 1 << 30
    x = torch.empty((n, 2), dtype=torch.float32, device=dev)
    piece = 1 << 27
    uctx = cb.Context(device=dev.index, cuda_stream=stream.cuda_stream)
    d_w = torch.empty(piece, dtype=torch.int32, device=dev)
    for o in range(0, n, piece):
        uctx.synth_words(d_w, piece, n0=o, tone_i=ti, tone_q=tq)
        uctx.unpack_submit(d_w, cf32=x[o:o + piece])
    uctx.wait()
    uctx.close()
    del d_w
    y = torch.zeros((n // DECIM, 2), dtype=torch.float32, device=dev)
    for order, fc, name in ((4, 250e3, "C3"), (6, 25e3, "C3_order6_fc25k")):
        ctx = cb.Context(device=dev.index, sos=cb.butter_lowpass(order, 4e6, fc), decimation=DECIM, parity=cb.PARITY_F32,
                         cuda_stream=stream.cuda_stream)
        ctx.set_profiling(True)
        mean_ms, best_ms = time_launches(torch, stream, lambda: ctx.filter_submit(cb.IN_CF32, x, n, y), lambda: ctx.wait(), reps=6, warm=2)
        km = ctx.read_profile()
        rec(name, "2^30 CF32 samples -> order-%d Butterworth fc = %g kHz -> decimate by 8, CF32 out" % (order, fc / 1e3),
            n, 8 + 8 / DECIM, mean_ms, best_ms, {"scan_kernel_ms": float(np.mean(km)) if len(km) else None})
        ctx.close()
    del x, y
    # ---- the reference's own Soapy configuration on the device: CS16 in place, order 6, full rate (8 B/sample)
    n = 1 << 28
    d_w = torch.empty(n, dtype=torch.int32, device=dev)
    cs = torch.zeros((n, 2), dtype=torch.int16, device=dev)
    uctx = cb.Context(device=dev.index, cuda_stream=stream.cuda_stream)
    uctx.synth_words(d_w, n, tone_i=ti, tone_q=tq)
    uctx.unpack_submit(d_w, cs16=cs)
    uctx.wait(); uctx.close()
    for fc, name in ((10e3, "soapy_q16_order6_fc10k"), (25e3, "soapy_q16_order6_fc25k"), (50e3, "soapy_q16_order6_fc50k")):
        ctx = cb.Context(device=dev.index, sos=cb.butter_lowpass(6, 4e6, fc), decimation=1, parity=cb.PARITY_Q16,
                         out_format=cb.FMT_CS16, cuda_stream=stream.cuda_stream)
        o16 = torch.zeros((n, 2), dtype=torch.int16, device=dev)
        mean_ms, best_ms = time_launches(torch, stream, lambda: ctx.filter_submit(cb.IN_CS16, cs, n, o16), lambda: ctx.wait(), reps=6, warm=2)
        rec(name, "2^28 CS16 samples -> order-6 fc = %g kHz at full rate -> int16 (CaribouliteStream.cpp:291-298)" % (fc / 1e3),
            n, 8.0, mean_ms, best_ms)
        ctx.close()
        del o16
    # words in, CF32 out at full rate (12 B/sample): the fused form of the same filters
    for order, fc, name in ((6, 10e3, "fused_d1_order6_fc10k"), (6, 50e3, "fused_d1_order6_fc50k"), (4, 250e3, "fused_d1_order4_fc250k")):
        ctx = cb.Context(device=dev.index, sos=cb.butter_lowpass(order, 4e6, fc), decimation=1, parity=cb.PARITY_F32,
                         cuda_stream=stream.cuda_stream)
        of = torch.zeros((n, 2), dtype=torch.float32, device=dev)
        mean_ms, best_ms = time_launches(torch, stream, lambda: ctx.filter_submit(cb.IN_WORDS, d_w, n, of), lambda: ctx.wait(), reps=6, warm=2)
        rec(name, "2^28 words -> order-%d fc = %g kHz at full rate -> CF32" % (order, fc / 1e3), n, 12.0, mean_ms, best_ms)
        ctx.close()
        del of
    del d_w, cs
    torch.cuda.empty_cache()

    # ---- e2e_shim: the calls a Soapy client makes, one MTU at a time, both sides on the same capture in a memfd
    out["e2e_shim"] = shim_leg(cb, torch, dev, with_cpu)
    if with_cpu:
        from oracle import pyoracle as po
        out["cpu_1thread"] = cpu_single_thread_legs(po)
    return out


def shim_leg(cb, torch, dev, with_cpu):
    """Stream::ReadSamplesGen(CF32, filter 50 kHz = type 2) per MTU on a memfd: caribou_cuda_rxstream_read (one queue:
    read(2) -> H2D -> unpack -> int16 filter -> convert -> D2H, one synchronisation) against the reference's own
    loop on one host thread.  The real device delivers one MTU every 32.8 ms."""
    L = cb.lib()
    n = 1 << 24
    ti, tq = cb.tone_tables()
    g = cb.Context(device=dev.index)
    d_w = torch.empty(n, dtype=torch.int32, device=dev)
    g.synth_words(d_w, n, tone_i=ti, tone_q=tq)
    g.wait(); g.close()
    raw = d_w.cpu().numpy().view(np.uint8)
    del d_w
    fd = os.memfd_create("bench_smi")
    os.write(fd, raw.tobytes())
    res = {"what": "Stream::ReadSamplesGen per MTU (%d samples) over a %d-MTU capture in a memfd; the filter is the Stream's "
                   "type 2 (order 6, fc 25 kHz)" % (MTU, n // MTU)}
    buf = np.zeros((MTU + 1, 2), dtype=np.float32)
    for fmt, filt, key in ((b"CF32", 2, "gpu_cf32_filter"), (b"CF32", 0, "gpu_cf32_nofilter"), (b"CS16", 0, "gpu_cs16_nofilter")):
        os.lseek(fd, 0, os.SEEK_SET)
        h = L.caribou_cuda_smi_open(fd, 524288, dev.index)
        st = L.caribou_cuda_rxstream_create(h, 0)
        L.caribou_cuda_rxstream_set_format(st, fmt)
        L.caribou_cuda_rxstream_set_filter(st, filt)
        for _ in range(4):                                    # warm-up reads (allocations, first launches)
            L.caribou_cuda_rxstream_read(st, buf.ctypes.data, MTU, 100000)
        got, lat = 0, []
        t0 = time.perf_counter()
        while True:
            t1 = time.perf_counter()
            r = L.caribou_cuda_rxstream_read(st, buf.ctypes.data, MTU, 100000)
            if r <= 0:
                break
            lat.append(time.perf_counter() - t1)
            got += r
        dt = time.perf_counter() - t0
        L.caribou_cuda_rxstream_destroy(st)
        L.caribou_cuda_smi_close(h)
        res[key] = {"msps": got / dt / 1e6, "ms_per_mtu_median": float(np.median(lat)) * 1e3,
                    "ms_per_mtu_p99": float(np.percentile(lat, 99)) * 1e3, "mtus": len(lat),
                    "radios_one_host_thread_serves": 32.768 / (float(np.median(lat)) * 1e3)}
    os.close(fd)
    if with_cpu:
        from oracle import pyoracle as po
        if po.ref_available():
            devr = po.RefDev(raw)
            devr.stream_open()
            devr.stream_set_format("CF32")
            devr.stream_set_filter(2)
            got = 0
            t0 = time.perf_counter()
            while True:
                r = devr.stream_read_into(buf, MTU)
                if r <= 0:
                    break
                got += r
            dt = time.perf_counter() - t0
            devr.close()
            res["reference_cf32_filter"] = {"msps": got / dt / 1e6, "ms_per_mtu": dt / max(got // MTU, 1) * 1e3, "threads": 1,
                                            "kind": "reference (oracle/_ref: unmodified CaribouliteStream.cpp + caribou_smi.c)"}
            res["ratio"] = res["gpu_cf32_filter"]["msps"] / res["reference_cf32_filter"]["msps"]
    return res


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
