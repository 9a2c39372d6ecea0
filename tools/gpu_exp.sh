# quick configuration experiments: device arm only
mkdir -p gpurun_out
Q="--steps 8 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 8"
for cfg in "4 592" "8 592" "2 592" "8 1184" "1 148" "1 296"; do
  set -- $cfg
  timeout 600 python bench.py $Q --handles $1 --lanes $2 > gpurun_out/exp_h$1_l$2.log 2> gpurun_out/exp_h$1_l$2.err || tail -3 gpurun_out/exp_h$1_l$2.err
  python - <<PY
import json
try:
    l=json.loads(open('gpurun_out/exp_h$1_l$2.log').read().strip().splitlines()[-1])
    print('handles $1 lanes $2: %.0f scans/s, %.2f ms/step, submit %s' % (l['value'], l['ms_per_step'], l['host_submit_ms_per_scan']))
except Exception as e: print('handles $1 lanes $2 failed', e)
PY
done
timeout 600 python bench.py $Q --handles 4 --lanes 592 --stagger 1 > gpurun_out/exp_stagger.log 2>/dev/null; python -c "
import json; l=json.loads(open('gpurun_out/exp_stagger.log').read().strip().splitlines()[-1]); print('stagger: %.0f' % l['value'])"
