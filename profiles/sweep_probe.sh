#!/bin/bash
# probe.py against every in-tree variant of the library; usage: sweep_probe.sh out.txt case [case ...]
out=$1; shift
: > $out
for lib in cariboulite_b200/libcaribou_cuda*.so; do
  echo "== $lib" >> $out
  CARIBOU_CUDA_LIB=$PWD/$lib timeout 300 python profiles/probe.py "$@" 2>&1 | python -c '
import sys, json
for ln in sys.stdin:
    try:
        j = json.loads(ln); print("%-18s kernel_ms %.4f  frac %.4f" % (j["case"], j["kernel_ms_avg"], j["frac"]))
    except Exception:
        print(ln.rstrip()[:200])' >> $out
done
cat $out
