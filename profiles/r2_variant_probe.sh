for v in "" _p3u5; do
 echo "== lib$v"
 CARIBOU_CUDA_LIB=$PWD/cariboulite_b200/libcaribou_cuda$v.so ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:iir_scan -s 2 -c 3 --csv python profiles/probe.py fused:64:24:4:8 2>/dev/null | grep -E "gpu__time|inst_exec" | awk -F'","' '{print $(NF-2), $NF}'
 CARIBOU_CUDA_LIB=$PWD/cariboulite_b200/libcaribou_cuda$v.so python profiles/probe.py fused:64:24:4:8 fused:64:24:6:8 fused:64:24:4:16 cf32:30:4:8 | python -c "
import json,sys
for l in sys.stdin:
    j=json.loads(l); print('  %-20s %.4f ms'%(j['case'],j['kernel_ms_avg']))"
done
