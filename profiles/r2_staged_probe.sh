CASES="fused:64:24:4:8 fused:64:24:2:8 fused:64:24:6:8 fused:64:24:6:8:10e3 fused:64:24:8:8 fused:64:24:4:16 fused:64:24:4:32 fused:64:24:4:4 fused:64:24:4:2 fused:64:24:6:4 cf32:30:4:8 cf32:30:6:8"
python -m pytest tests/test_gpu_iir.py tests/test_gpu_fullsize.py tests/test_gpu_shard.py tests/test_gpu_batch.py tests/test_gpu_compat.py -m gpu -q -x > gpurun_out/r2f_tests.log 2>&1; tail -5 gpurun_out/r2f_tests.log
rm -f gpurun_out/r2f_probe.txt
for v in ${VARIANTS:-main _defer}; do v=${v#main}
  echo "== lib$v" >> gpurun_out/r2f_probe.txt; CARIBOU_CUDA_LIB=$PWD/cariboulite_b200/libcaribou_cuda$v.so python profiles/probe.py $CASES >> gpurun_out/r2f_probe.txt 2>&1
done
python - <<'PY'
import json
for l in open('gpurun_out/r2f_probe.txt'):
    if l.startswith('=='): print(l.strip()); continue
    try: j=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print("%-24s kernel %.3f ms  frac %.3f  %6.1f Gsps run_len %s window %s"%(j['case'],j['kernel_ms_avg'],j['frac'],j['Gsps'],j.get('run_len'),j.get('window')))
PY
