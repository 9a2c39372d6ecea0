mkdir -p gpurun_out
Q="--steps 6 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 8 --handles 1 --lanes 592"
for K in 12 16; do
  LEGO_NVCC_DEFS="-DLEGO_KNN_K=$K" python -c "
import lego_loam_b200
from lego_loam_b200.build import build
build(force=True)" > gpurun_out/k$K.build.log 2>&1 || { tail -5 gpurun_out/k$K.build.log; continue; }
  timeout 600 python bench.py $Q --profile-json gpurun_out/k${K}_prof.json > gpurun_out/k$K.log 2> gpurun_out/k$K.err || tail -3 gpurun_out/k$K.err
  python -c "
import json
p=json.load(open('gpurun_out/k${K}_prof.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
print('K=$K value', round(p['line']['value']), 'sum', round(sum(v['ms'] for v in k0.values()),1), {k: round(v['ms'],2) for k,v in k0.items() if k.startswith('mo_')})"
done
