#!/bin/bash
# runs bench.py's device-resident leg against every tuning variant of the library present in-tree
# (built with CBL_VARIANT=<tag> CBL_NVCC_EXTRA=... python -m cariboulite_b200.build); one line per variant
out=${1:-gpurun_out/sweep.txt}
: > $out
for lib in cariboulite_b200/libcaribou_cuda*.so; do
  r=$(CARIBOU_CUDA_LIB=$PWD/$lib timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e 2>&1 | tail -1)
  echo "$lib $(echo "$r" | python -c 'import sys,json
try:
    j=json.loads(sys.stdin.read()); r=j["roofline"]; print("kernel_ms_avg=%.4f min=%.4f frac=%.4f ms_per_step=%.4f" % (r["kernel_ms_avg"], r["kernel_ms_min"], r["frac"], j["ms_per_step"]))
except Exception as e:
    print("FAILED", e)')" >> $out
done
cat $out
