#!/usr/bin/env python
"""C5 in small on real GPUs (run under torchrun, one rank per GPU, NCCL):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tests/multi_gpu_c5.py [--log2-words 24] [--mode allgather|chain] [--calls 2]

One continuous stream of G * calls * 2^log2 words, device-generated in place, time-chunked over the G ranks;
fused unpack -> order-6 Butterworth (fc = 10 kHz, the narrowest reference filter) -> decimate by 8 with the
IIR state handed from GPU to GPU ON THE DEVICE: caribou_cuda_shard_submit (ncclAllGather or the ncclSend/ncclRecv
chain on the context's stream, composition kernel, fix-up from device memory; the communicator is created inside
the library from a unique id that rank 0 ships through torch.distributed).  Rank 0 gathers every rank's output and compares the WHOLE stream with the
sequential CPU oracle (tolerance: 1e-5 of signal RMS per rail, north_star); it also prints the time of the
hand-off.  Prints one JSON line.  tests/test_gpu_multi.py runs it when the box has >= 2 GPUs."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2-words", type=int, default=22)
    ap.add_argument("--mode", default="allgather")
    ap.add_argument("--calls", type=int, default=2)
    ap.add_argument("--order", type=int, default=6)
    ap.add_argument("--fc", type=float, default=10e3)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import cariboulite_b200 as cb
    from oracle import pyoracle as po

    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    n, D = 1 << a.log2_words, 8
    sos = cb.butter_lowpass(a.order, 4e6, a.fc)
    ti, tq = cb.tone_tables()
    ctx = cb.Context(device=local, sos=sos, decimation=D, parity=cb.PARITY_F32, out_format=cb.FMT_CF32)
    ids = [cb.shard_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    ctx.shard_init(rank, world, ids[0])
    default_mode = ctx.shard_mode
    if a.mode != "default":
        ctx.shard_set_mode({"chain": cb.SHARD_CHAIN, "allgather": cb.SHARD_ALLGATHER, "peer": cb.SHARD_PEER}[a.mode])
    ctx.set_profiling(True)
    d_w = torch.empty(n, dtype=torch.int32, device=dev)
    d_out = torch.zeros((a.calls, n // D, 2), dtype=torch.float32, device=dev)
    t_hand, t_dev = [], []
    for call in range(a.calls):
        n0 = (call * world + rank) * n
        ctx.synth_words(d_w, n, n0=n0, tone_i=ti, tone_q=tq)
        torch.cuda.synchronize(dev)
        dist.barrier()
        t0 = time.perf_counter()
        ctx.shard_submit(cb.IN_WORDS, d_w, n, d_out[call])
        assert ctx.wait()[0] == n
        torch.cuda.synchronize(dev)
        t_hand.append(time.perf_counter() - t0)
        t_dev.append(ctx.shard_profile())
    # gather on rank 0: [world][calls][n/D][2]
    allout = [torch.empty_like(d_out) for _ in range(world)] if rank == 0 else None
    dist.gather(d_out, allout, dst=0)
    ok, err = True, None
    if rank == 0:
        total = a.calls * world * n
        w = po.gen_words(total)
        ref = po.rx_path_f32(w, po.butter_lowpass_sos(a.order, 4e6, a.fc), D)
        got = np.zeros_like(ref)
        for r in range(world):
            o = allout[r].cpu().numpy()
            for call in range(a.calls):
                lo = (call * world + r) * n // D
                got[lo:lo + n // D] = o[call]
        e = np.sqrt(np.mean((got.astype(np.float64) - ref) ** 2, axis=0)) / np.sqrt(np.mean(ref.astype(np.float64) ** 2, axis=0))
        # windows straddling every shard boundary, separately (a hand-off bug hides in a whole-stream RMS)
        wmax = 0.0
        for b in range(1, a.calls * world):
            c = b * n // D
            seg = slice(c - 64, c + 4096)
            ew = np.sqrt(np.mean((got[seg].astype(np.float64) - ref[seg]) ** 2)) / np.sqrt(np.mean(ref[seg].astype(np.float64) ** 2))
            wmax = max(wmax, float(ew))
        err = {"rel_rms": [float(v) for v in e], "worst_boundary_window": wmax}
        ok = bool(np.all(e <= 1e-5) and wmax <= 1e-5)
        print(json.dumps({"world": world, "mode": a.mode, "default_mode": default_mode, "mode_used": ctx.shard_mode, "words_per_rank_per_call": n, "calls": a.calls, "ok": ok,
                          "err": err, "fixup_samples": int(ctx.fixup_samples(n)),
                          "process_ms_per_call": [round(t * 1e3, 3) for t in t_hand],
                          "pass_ms": [round(t[0], 4) for t in t_dev], "handoff_fixup_ms": [round(t[1], 4) for t in t_dev]}),
              flush=True)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, src=0)
    ctx.close()
    dist.destroy_process_group()
    return 0 if int(flag.item()) == 1 else 1


if __name__ == "__main__":
    sys.exit(main())
