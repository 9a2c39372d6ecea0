mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_map_optimization.py -x -q -m gpu -k "incremental" > gpurun_out/r2_incr.log 2>&1; echo "incr rc=$?"; tail -15 gpurun_out/r2_incr.log
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_pytest.log
Q="--steps 8 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 8"
timeout 600 python bench.py $Q --handles 4 --lanes 592 > gpurun_out/r2_h4.log 2> gpurun_out/r2_h4.err || tail -5 gpurun_out/r2_h4.err
timeout 600 python bench.py $Q --handles 1 --lanes 592 --profile-json gpurun_out/r2_prof_h1_l592.json > gpurun_out/r2_h1.log 2> gpurun_out/r2_h1.err || tail -5 gpurun_out/r2_h1.err
python -c "
import json
for f in ('r2_h4','r2_h1'):
    l=json.loads(open('gpurun_out/%s.log'%f).read().strip().splitlines()[-1]); print(f, round(l['value']), l['config']['local_map'])"
