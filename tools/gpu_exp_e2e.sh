# arena uploads: API tests, then the e2e arm with one copy per lane (0) vs runs of lanes (3), default window
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pipeline_api.py tests/test_gpu_map_optimization.py -x -q -m gpu 2>&1 | tail -4
Q="--steps 8 --warmup 5 --no-cpu-baseline --no-latency --secondary none"
for C in 0 3; do
  timeout 900 python bench.py $Q --host-copies $C > gpurun_out/e2e_c$C.log 2> gpurun_out/e2e_c$C.err || tail -3 gpurun_out/e2e_c$C.err
  python -c "
import json
l=json.loads(open('gpurun_out/e2e_c$C.log').read().strip().splitlines()[-1])
print('copies=$C value', round(l['value']), 'e2e', round(l['e2e']['value']), {k: l['e2e'][k] for k in ('h2d_bytes_per_scan','h2d_gbs_achieved','pcie_h2d_gbs_plain_copy','poses_equal_device_arm') if k in l['e2e']}, l['e2e'].get('h2d_copies_per_scan'), 'launches', l.get('gpu_launches'))"
done
