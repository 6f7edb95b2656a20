# final scaling evidence on one 8-GPU box: bench.py at N = 2, 4, 8 (C4 value, e2e with ceiling, c5 with hand-off), the multi-GPU tests
set -x
nvidia-smi topo -m > gpurun_out/r2_topo8.txt 2>&1
for N in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N 2> gpurun_out/r2_scale_${N}.err | grep "^{" | tail -1 > gpurun_out/r2_scale_${N}gpu.json
done
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_batch.py::test_contexts_on_two_devices_in_one_process -m gpu -q > gpurun_out/r2_multi_tests.log 2>&1; tail -3 gpurun_out/r2_multi_tests.log
python - <<'PY'
import json
for N in (2,4,8):
    try:
        d=json.loads(open("gpurun_out/r2_scale_%dgpu.json"%N).read())
    except Exception as e:
        print(N,"no line",e); continue
    c=d.get("c5") or {}
    print(N, "value %.0f  e2e %.0f (ceil %.0f)  c5 %.0f ms/call %.3f handoff %.1f/%.1f us mode %s rel %.2g/%.2g eff %.3f"%(d["value"], d["e2e"]["value"], d["e2e"].get("ceiling_msps",0), c.get("value",0), c.get("ms_per_call",0), 1e3*c.get("handoff_ms",0), 1e3*c.get("handoff_ms_min_over_ranks",0), c.get("handoff_mode"), c.get("rel_rms",0), c.get("worst_boundary_window",0), c.get("efficiency_vs_c4_kernel",0)))
PY
