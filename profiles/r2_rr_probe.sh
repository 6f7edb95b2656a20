CASES="fused:1:30:2:1 fused:1:30:4:1 fused:1:30:6:1 fused:1:30:6:1:10e3 fused:1:30:6:1:50e3 fused:1:30:8:1 fused:1:30:6:2 fused:1:30:8:2 cf32:28:4:1 cf32:28:6:1"
python -m pytest tests/test_gpu_rerun.py tests/test_gpu_iir.py tests/test_gpu_shard.py tests/test_gpu_compat.py tests/test_gpu_batch.py -m gpu -q -x > gpurun_out/r2d_tests.log 2>&1; tail -5 gpurun_out/r2d_tests.log
rm -f gpurun_out/r2d_probe.txt
for v in "" _w16r3 _p3u2 _lt512 _lt128; do
  echo "== lib$v" >> gpurun_out/r2d_probe.txt; CARIBOU_CUDA_LIB=$PWD/cariboulite_b200/libcaribou_cuda$v.so python profiles/probe.py $CASES >> gpurun_out/r2d_probe.txt 2>&1
done
python - <<'PY'
import json
for l in open('gpurun_out/r2d_probe.txt'):
    if l.startswith('=='): print(l.strip()); continue
    try: j=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print("%-24s kernel %.3f ms  frac %.3f  %6.1f Gsps run_len %s window %s"%(j['case'],j['kernel_ms_avg'],j['frac'],j['Gsps'],j.get('run_len'),j.get('window')))
PY
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum
for c in fused:1:30:2:1 fused:1:30:6:1:10e3; do
  echo "== ncu main $c"
  ncu --metrics $M --clock-control none -k regex:iir_rerun -c 2 --csv python profiles/probe.py $c 2>/dev/null | grep -v "^==" | python -c "
import csv,sys
rows=[r for r in csv.reader(sys.stdin) if len(r)>10]
h=rows[0]; i_id=h.index('ID'); i_m=h.index('Metric Name'); i_v=h.index('Metric Value')
last=max(int(r[i_id]) for r in rows[1:])
print('   '+'  '.join('%s=%s'%(r[i_m].split('.')[0].replace('__','_')[-22:],r[i_v]) for r in rows[1:] if int(r[i_id])==last))
"
done
