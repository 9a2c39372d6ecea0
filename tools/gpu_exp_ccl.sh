# tiled connected components + staged fa_prepare: parity tests, then per-kernel profile (592 lanes, one handle) old vs new
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_image_projection.py tests/test_gpu_feature_association.py tests/test_gpu_quirks.py tests/test_gpu_golden.py -x -q -m gpu 2>&1 | tail -6
Q="--steps 4 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none"
for T in 0 1; do
LEGO_CCL_TILE=$T timeout 600 python bench.py $Q --handles 1 --lanes 592 --profile-json gpurun_out/ccl$T.json > gpurun_out/ccl$T.log 2> gpurun_out/ccl$T.err || tail -3 gpurun_out/ccl$T.err
python -c "
import json
p=json.load(open('gpurun_out/ccl$T.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
tot=sum(v['ms'] for v in k0.values()); print('TILE=$T value', round(p['line']['value']), 'sum ms', round(tot,2))
for k,v in sorted(k0.items(), key=lambda kv:-kv[1]['ms']):
    if k.startswith('ip_') or k.startswith('fa_prep') or k.startswith('memset'): print('   %-32s %7.3f ms %5.1f%% n=%s'%(k,v['ms'],100*v['ms']/tot,v.get('launches')))"
done
LEGO_CCL_TILE=1 timeout 600 python bench.py --steps 8 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none > gpurun_out/ccl_tp.log 2> gpurun_out/ccl_tp.err; python -c "
import json
l=json.loads(open('gpurun_out/ccl_tp.log').read().strip().splitlines()[-1]); print('592x4 value', round(l['value']), 'ms/step', round(l['ms_per_step'],2))"
