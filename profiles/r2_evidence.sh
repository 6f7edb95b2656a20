# Round-2 evidence run on one B200 (the numbers under profiles/r2_* come from this script):
#   GPU test suite, bench line, reference line, launch list of the bench (cold-cache, serialised: shares only),
#   ncu --set full captures of the dominant kernel of each family.
set -x
python -m pytest tests -m gpu -q > gpurun_out/r2_gputests.log 2>&1; tail -3 gpurun_out/r2_gputests.log
python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench_line.json 2> gpurun_out/r2_bench.err; tail -c 400 gpurun_out/r2_bench_line.json
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2_reference_line.json 2> gpurun_out/r2_ref.err; tail -c 600 gpurun_out/r2_reference_line.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-configs --no-sustain --no-c5 > gpurun_out/r2_ncu_bench.log 2>&1
FULL="ncu --set full --clock-control none --import-source on"
$FULL -k regex:iir_scan_kernel -s 3 -c 1 -f -o gpurun_out/r2_scan_c4 python profiles/probe.py fused:64:24:4:8 > /dev/null 2>&1
$FULL -k regex:iir_rerun_kernel -s 3 -c 1 -f -o gpurun_out/r2_rr_6_1_10e3 python profiles/probe.py fused:1:30:6:1:10e3 > /dev/null 2>&1
$FULL -k regex:iir_rerun_kernel -s 3 -c 1 -f -o gpurun_out/r2_rr_4_1 python profiles/probe.py fused:1:30:4:1 > /dev/null 2>&1
$FULL -k regex:unpack_kernel -s 3 -c 1 -f -o gpurun_out/r2_unpack python profiles/probe.py unpack:26:0 > /dev/null 2>&1
$FULL -k regex:tx_pack_kernel -s 3 -c 1 -f -o gpurun_out/r2_tx python profiles/probe.py tx:28:CS16 > /dev/null 2>&1
$FULL -k regex:find_offsets_kernel -s 3 -c 1 -f -o gpurun_out/r2_find python profiles/probe.py unpack:26:0 > /dev/null 2>&1
python profiles/probe.py fused:64:24:4:8 fused:64:24:2:8 fused:64:24:6:8 fused:64:24:6:8:10e3 fused:64:24:8:8 fused:64:24:4:16 fused:64:24:4:32 fused:64:24:4:4 fused:64:24:4:2 fused:1:30:2:1 fused:1:30:4:1 fused:1:30:6:1 fused:1:30:6:1:10e3 fused:1:30:6:1:50e3 fused:1:30:8:1 fused:1:30:6:2 fused:1:30:8:2 cf32:30:4:8 cf32:30:6:8 cf32:28:4:1 cf32:28:6:1 unpack:26:0 unpack:26:1 unpack:26:131072 unpack:22:0 tx:28:CS16 tx:28:CF32 tx:27:CF64 tx:28:CS8 > gpurun_out/r2_probes.txt 2>&1
ls -la gpurun_out/*.ncu-rep | tail
