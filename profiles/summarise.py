#!/usr/bin/env python
"""Summarises an ncu report (read here, without a GPU) into the text files committed under profiles/.

    python profiles/summarise.py gpurun_out/r1_iir.ncu-rep profiles/r1_iir  [--captures 64 --log2-words 24]

Writes <out>_raw.txt (the handful of raw metrics the roofline uses), <out>_sass_mix.txt (instruction mix and
the instructions that collect most stall samples, from the source page; needs -lineinfo) and, with
--captures/--log2-words, profiles/roofline_traffic.json (dram bytes per launch, read by bench.py)."""
import argparse, collections, csv, io, json, os, subprocess, sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__cycles_elapsed.max", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def ncu_csv(rep, page):
    r = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True)
    return r.stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep"); ap.add_argument("out")
    ap.add_argument("--captures", type=int); ap.add_argument("--log2-words", type=int)
    ap.add_argument("--samples", type=float, default=0, help="complex samples per launch (for per-sample figures)")
    a = ap.parse_args()
    rows = list(csv.reader(io.StringIO(ncu_csv(a.rep, "raw"))))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(a.out + "_raw.txt", "w") as f:
        f.write("# ncu --set full --clock-control none, %s\n" % os.path.basename(a.rep))
        ki = hdr.index("Kernel Name")
        for r in data:
            f.write("# launch: %s\n" % r[ki][:160])
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                f.write("%-90s %-12s %s\n" % (k, units[i], "  ".join(r[i] for r in data)))
    gb = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "Tbyte": 1e12}
    tot = []
    for r in data:
        t = 0.0
        for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = hdr.index(k)
            t += float(r[i]) * gb[units[i]]
        tot.append(t)
    if a.captures:
        with open(os.path.join(os.path.dirname(a.out) or ".", "roofline_traffic.json"), "w") as f:
            json.dump({"captures": a.captures, "log2_words": a.log2_words, "dram_bytes_per_launch": sum(tot) / len(tot),
                       "per_launch": tot, "source": os.path.basename(a.rep)}, f, indent=1)
    # ---- source page: first kernel only
    src = ncu_csv(a.rep, "source")
    blocks = src.split('"Kernel Name"')
    if len(blocks) < 2:
        return
    rows = list(csv.reader(io.StringIO(blocks[1].split("\n", 1)[1])))
    h = rows[0]
    iS, iN, iE = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
    stall_cols = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    mix = collections.Counter(); samp = collections.Counter(); total_e = total_s = 0
    inst = []
    for r in rows[1:]:
        if len(r) <= max(iS, iN, iE) or not r[iE].isdigit():
            continue
        op = r[iS].strip().split()
        if op and op[0].startswith("@"):
            op = op[1:]
        opc = op[0].rstrip(";") if op else "?"
        e, s = int(r[iE]), int(r[iN])
        mix[opc] += e; samp[opc] += s; total_e += e; total_s += s
        st = sorted(((int(r[i]) if r[i].isdigit() else 0, h[i]) for i in stall_cols), reverse=True)[:2]
        inst.append((s, e, r[iS].strip()[:70], st))
    with open(a.out + "_sass_mix.txt", "w") as f:
        per = (a.samples / 32.0) if a.samples else None
        f.write("# warp instructions executed by opcode (first profiled launch); stall samples by opcode\n")
        f.write("# total warp instructions %d, total stall samples %d%s\n" %
                (total_e, total_s, (", %.2f warp instructions per sample per lane" % (total_e / per)) if per else ""))
        for opc, e in mix.most_common(30):
            f.write("%-16s %12d %6.2f%%  samples %6.2f%%%s\n" % (opc, e, 100.0 * e / total_e, 100.0 * samp[opc] / max(1, total_s),
                                                               ("  %.3f/sample" % (e / per)) if per else ""))
        f.write("\n# 40 instructions with most stall samples: samples, executed, SASS, top stall reasons\n")
        for s, e, txt, st in sorted(inst, reverse=True)[:40]:
            f.write("%7d %10d  %-70s %s\n" % (s, e, txt, " ".join("%s=%d" % (n, v) for v, n in st if v)))


if __name__ == "__main__":
    main()
