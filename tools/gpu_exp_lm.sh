# persistent per-lane LM kernel: parity A/B test, then block size / register budget variants against the launch loop
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_map_optimization.py -x -q -m gpu 2>&1 | tail -5
Q="--steps 6 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 8 --lanes 592"
run() {  # name, env, handles
  env $2 timeout 600 python bench.py $Q --handles $3 --profile-json gpurun_out/lm_$1_h$3.json > gpurun_out/lm_$1_h$3.log 2> gpurun_out/lm_$1_h$3.err || tail -3 gpurun_out/lm_$1_h$3.err
  python -c "
import json
p=json.load(open('gpurun_out/lm_$1_h$3.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
print('$1 h$3 value', round(p['line']['value']), 'sum', round(sum(v['ms'] for v in k0.values()),1), {k: round(v['ms'],2) for k,v in k0.items() if k.startswith('mo_')})"
}
run loop LEGO_LM_PERSISTENT=0 1
run loop LEGO_LM_PERSISTENT=0 4
run t512b2 LEGO_LM_PERSISTENT=1 1
run t512b2 LEGO_LM_PERSISTENT=1 4
for V in "256 2" "256 4" "512 1" "1024 1"; do
  set -- $V
  LEGO_NVCC_DEFS="-DLEGO_MO_LM_THREADS=$1 -DLEGO_MO_LM_MINBLOCKS=$2" python -c "
import lego_loam_b200
from lego_loam_b200.build import build
build(force=True)" > gpurun_out/lm_build_$1_$2.log 2>&1 || { tail -5 gpurun_out/lm_build_$1_$2.log; continue; }
  run t$1b$2 LEGO_LM_PERSISTENT=1 1
  run t$1b$2 LEGO_LM_PERSISTENT=1 4
done
