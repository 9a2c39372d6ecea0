# image projection: loads hoisted out of the dependent chains (ccl runs / union / stats, compaction flags)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_image_projection.py tests/test_gpu_quirks.py tests/test_gpu_stage_api.py -x -q -m gpu 2>&1 | tail -4
Q="--steps 6 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 8 --lanes 592"
for H in 1 4; do
  timeout 600 python bench.py $Q --handles $H --profile-json gpurun_out/ip_h$H.json > gpurun_out/ip_h$H.log 2> gpurun_out/ip_h$H.err || tail -3 gpurun_out/ip_h$H.err
  python -c "
import json
p=json.load(open('gpurun_out/ip_h$H.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
print('h$H value', round(p['line']['value']), 'sum', round(sum(v['ms'] for v in k0.values()),1), {k: round(v['ms'],2) for k,v in k0.items() if k.startswith('ip_')})"
done
