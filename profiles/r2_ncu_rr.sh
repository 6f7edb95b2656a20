# ncu metrics of the rerun kernel for a few filter orders (tuning aid)
M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__inst_executed.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_op_read.sum
for c in fused:1:30:2:1 fused:1:30:6:1:10e3; do
  ncu --metrics $M --clock-control none -k regex:iir_rerun -c 3 --csv python profiles/probe.py $c 2>/dev/null | grep -v "^==" > gpurun_out/r2_ncu_$(echo $c | tr ':' '_').csv
done
python - <<'PY'
import csv,glob
for f in sorted(glob.glob('gpurun_out/r2_ncu_fused*.csv')):
    rows=[r for r in csv.reader(open(f)) if len(r)>10]
    if not rows: print(f,"empty"); continue
    h=rows[0]; i_id=h.index('ID'); i_m=h.index('Metric Name'); i_v=h.index('Metric Value')
    last=max(int(r[i_id]) for r in rows[1:])
    print("==",f)
    for r in rows[1:]:
        if int(r[i_id])==last: print("  %-90s %s"%(r[i_m],r[i_v]))
PY
