M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active
for v in "" _lt256nh _lt128 _lt128nh _lt64; do
 for c in fused:1:30:2:1 fused:1:30:6:1:10e3; do
  echo "== lib$v $c"
  CARIBOU_CUDA_LIB=$PWD/cariboulite_b200/libcaribou_cuda$v.so ncu --metrics $M --clock-control none -k regex:iir_rerun -c 2 --csv python profiles/probe.py $c 2>/dev/null | grep -v "^==" | python -c "
import csv,sys
rows=[r for r in csv.reader(sys.stdin) if len(r)>10]
h=rows[0]; i_id=h.index('ID'); i_m=h.index('Metric Name'); i_v=h.index('Metric Value')
last=max(int(r[i_id]) for r in rows[1:])
print('   '+'  '.join('%s=%s'%(r[i_m].split('.')[0].replace('__','_')[-22:],r[i_v]) for r in rows[1:] if int(r[i_id])==last))
"
 done
done
