mkdir -p gpurun_out
Q="--steps 6 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 8"
for T in 256 384; do
  LEGO_NVCC_DEFS="-DLEGO_ODO_THREADS=$T" python -c "
import lego_loam_b200
from lego_loam_b200.build import build
build(force=True)" > gpurun_out/odo$T.build.log 2>&1 || { tail -5 gpurun_out/odo$T.build.log; continue; }
  timeout 600 python bench.py $Q --handles 1 --lanes 592 --profile-json gpurun_out/odo${T}_prof.json > gpurun_out/odo$T.log 2> gpurun_out/odo$T.err || tail -3 gpurun_out/odo$T.err
  timeout 600 python bench.py $Q --handles 4 --lanes 592 > gpurun_out/odo${T}_h4.log 2> gpurun_out/odo${T}_h4.err || tail -3 gpurun_out/odo${T}_h4.err
  timeout 300 python -m pytest tests/test_gpu_feature_association.py -x -q -m gpu -k "vlp16-street or hdl64" 2>&1 | tail -1
  python -c "
import json
p=json.load(open('gpurun_out/odo${T}_prof.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
l4=json.loads(open('gpurun_out/odo${T}_h4.log').read().strip().splitlines()[-1])
print('ODO_THREADS=$T h1', round(p['line']['value']), 'h4', round(l4['value']), 'sum', round(sum(v['ms'] for v in k0.values()),1), 'fa_odometry', round(k0['fa_odometry_kernel']['ms'],2))"
done
