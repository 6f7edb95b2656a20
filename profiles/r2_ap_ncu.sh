M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,sm__cycles_active.avg,smsp__inst_executed_op_dfma_pred_on.sum,sm__warps_active.avg.pct_of_peak_sustained_active
for env in "X=1" "CARIBOU_CUDA_NO_AP=1"; do
for c in fused:1:30:6:1:10e3; do
  echo "== ncu $env $c"
  env $env ncu --metrics $M --clock-control none -k regex:iir_rerun -c 2 --csv python profiles/probe.py $c 2>/dev/null | grep -v "^==" | python -c "
import csv,sys
rows=[r for r in csv.reader(sys.stdin) if len(r)>10]
h=rows[0]; i_id=h.index('ID'); i_m=h.index('Metric Name'); i_v=h.index('Metric Value')
last=max(int(r[i_id]) for r in rows[1:])
for r in rows[1:]:
    if int(r[i_id])==last: print('   %s = %s'%(r[i_m],r[i_v]))
"
done
done
