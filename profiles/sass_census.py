#!/usr/bin/env python
"""Opcode census of the shipped library (cuobjdump -sass libcaribou_cuda.so): per kernel the counts of the
instructions that identify how data moves and what computes - UTMALDG (TMA tensor loads), SYNCS (mbarrier),
STG/LDG by width, LDS/STS, DFMA/FFMA, F2F, SHFL, ATOM - so that the claims in DESIGN.md can be checked against the
object code.  Usage: python profiles/sass_census.py [lib.so] > profiles/r2_sass_census.txt"""
import collections, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "cariboulite_b200", "libcaribou_cuda.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
COLS = ["insts", "UTMALDG", "SYNCS", "DFMA", "FFMA", "F2F", "SHFL", "LDS.128", "LDS", "STS", "LDG.128", "LDG", "STG.256", "STG.128", "STG", "ATOM", "LDC+LDCU"]
per = collections.OrderedDict()
cur = None
for ln in txt.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(.*", "", name).replace("void ", "").replace("cbl::", "").replace("(int)", "").replace("(bool)", "")
        cur = per.setdefault(name, collections.Counter())
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", ln)
    if not m or cur is None:
        continue
    op = m.group(1)
    base = op.split(".")[0]
    cur["insts"] += 1
    if base == "UTMALDG": cur["UTMALDG"] += 1
    elif base == "SYNCS": cur["SYNCS"] += 1
    elif base in ("DFMA", "FFMA", "F2F", "SHFL", "STS"): cur[base] += 1
    elif base == "LDS": cur["LDS.128" if ".128" in op else "LDS"] += 1
    elif base == "LDG": cur["LDG.128" if ".128" in op else "LDG"] += 1
    elif base == "STG": cur["STG.256" if ".256" in op else ("STG.128" if ".128" in op else "STG")] += 1
    elif base in ("ATOM", "ATOMG", "RED"): cur["ATOM"] += 1
    elif base in ("LDC", "LDCU", "ULDC"): cur["LDC+LDCU"] += 1
print("# cuobjdump -sass %s (sm_100a), %d kernels" % (os.path.basename(lib), len(per)))
print("%-58s" % "kernel" + "".join("%9s" % c for c in COLS))
fam = collections.OrderedDict()
for name, c in per.items():
    f = re.sub(r"<.*", "", name)
    agg = fam.setdefault(f, [0, collections.Counter()])
    agg[0] += 1
    agg[1].update(c)
SHOW = ["iir_scan_kernel<2, 8, 0, true, true>", "iir_scan_kernel<3, 8, 0, true, true>", "iir_scan_kernel<2, 8, 2, true, true>",
        "iir_rerun_kernel<3, 1, 0>", "iir_rerun_kernel<3, 1, 1>", "iir_rerun_kernel<2, 1, 0>"]
for name, c in per.items():
    if name in SHOW or not name.startswith(("iir_scan_kernel", "iir_rerun_kernel", "iir_tail_kernel", "tx_pack_kernel")):
        print("%-58s" % name[:58] + "".join("%9d" % c[k] for k in COLS))
print("# families (all instantiations summed)")
for f, (n, c) in fam.items():
    print("%-58s" % ("%s  x%d" % (f, n)) + "".join("%9d" % c[k] for k in COLS))
