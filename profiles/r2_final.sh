# Final round-2 evidence on one box (2 GPUs so that the NCCL / peer tests run): GPU test suite, smoke, bench line,
# reference line, launch list of the bench (cold-cache, serialised: shares only), probes, ncu captures of the kernels
# that changed last (find_offsets, tail, order-4 AP rerun kernel)
set -x
python -m pytest tests -m gpu -q -x > gpurun_out/r2_gputests.log 2>&1; tail -3 gpurun_out/r2_gputests.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_smoke.log 2>&1; tail -2 gpurun_out/r2_smoke.log
python bench.py --steps 20 --warmup 3 2> gpurun_out/r2_bench.err | grep "^{" | tail -1 > gpurun_out/r2_bench_line.json; tail -c 300 gpurun_out/r2_bench_line.json
python bench.py --impl reference --steps 5 --warmup 1 2> gpurun_out/r2_ref.err | grep "^{" | tail -1 > gpurun_out/r2_reference_line.json; tail -c 400 gpurun_out/r2_reference_line.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-configs --no-sustain --no-c5 > gpurun_out/r2_ncu_bench.log 2>&1
python profiles/probe.py fused:64:24:4:8 fused:64:24:2:8 fused:64:24:6:8 fused:64:24:6:8:10e3 fused:64:24:8:8 fused:64:24:4:16 fused:64:24:4:32 fused:64:24:4:4 fused:64:24:4:2 fused:1:30:2:1 fused:1:30:4:1 fused:1:30:6:1 fused:1:30:6:1:10e3 fused:1:30:6:1:50e3 fused:1:30:8:1 fused:1:30:6:2 fused:1:30:8:2 cf32:30:4:8 cf32:30:6:8 cf32:28:4:1 cf32:28:6:1 unpack:26:0 unpack:26:1 unpack:26:131072 unpack:22:0 tx:28:CS16 tx:28:CF32 tx:27:CF64 tx:28:CS8 > gpurun_out/r2_probes.txt 2>&1
FULL="ncu --set full --clock-control none --import-source on"
$FULL -k regex:iir_rerun_kernel -s 3 -c 1 -f -o gpurun_out/r2_ap_4_1 python profiles/probe.py fused:1:30:4:1 > /dev/null 2>&1
$FULL -k regex:find_offsets_kernel -s 3 -c 1 -f -o gpurun_out/r2_find python profiles/probe.py unpack:26:0 > /dev/null 2>&1
ls -la gpurun_out/*.ncu-rep | tail -4
