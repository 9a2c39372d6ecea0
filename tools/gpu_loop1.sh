mkdir -p gpurun_out
timeout 900 python bench.py --workload loop --lanes 8 --handles 1 --distinct 8 --steps 11 --warmup 5 --no-latency --secondary none > gpurun_out/loop_l8.log 2> gpurun_out/loop_l8.err; echo "rc=$?"; tail -c 1500 gpurun_out/loop_l8.err
python -c "
import json
l=json.loads(open('gpurun_out/loop_l8.log').read().strip().splitlines()[-1])
print(round(l['value']), 'e2e', round(l['e2e']['value']), 'cpu', round(l['cpu_baseline']['value'],1), l['config']['timed_scan_range'], l['config']['local_map'])
print(l['cpu_baseline']['per_stage_ms'])"
