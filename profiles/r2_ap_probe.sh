# A/B of the rerun kernel's phase 3: AP form (integer FIR + all-pole cascade) against the parallel form (CARIBOU_CUDA_NO_AP=1)
CASES="fused:1:30:2:1 fused:1:30:4:1 fused:1:30:6:1 fused:1:30:6:1:10e3 fused:1:30:6:1:50e3 fused:1:30:8:1 fused:1:30:6:2 fused:1:30:8:2"
python -m pytest tests/test_gpu_rerun.py tests/test_gpu_iir.py tests/test_gpu_shard.py tests/test_gpu_compat.py tests/test_gpu_batch.py tests/test_gpu_edges.py tests/test_dropin.py -m gpu -q -x > gpurun_out/r2_ap_tests.log 2>&1; tail -15 gpurun_out/r2_ap_tests.log
rm -f gpurun_out/r2_ap_probe.txt
echo "== AP" >> gpurun_out/r2_ap_probe.txt; python profiles/probe.py $CASES >> gpurun_out/r2_ap_probe.txt 2>&1
echo "== parallel form (CARIBOU_CUDA_NO_AP=1)" >> gpurun_out/r2_ap_probe.txt; CARIBOU_CUDA_NO_AP=1 python profiles/probe.py $CASES >> gpurun_out/r2_ap_probe.txt 2>&1
python - <<'PY'
import json
for l in open('gpurun_out/r2_ap_probe.txt'):
    if l.startswith('=='): print(l.strip()); continue
    try: j=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print("%-24s kernel %.3f ms  frac %.3f  %6.1f Gsps run_len %s window %s"%(j['case'],j['kernel_ms_avg'],j['frac'],j['Gsps'],j.get('run_len'),j.get('window')))
PY
