mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r4_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r4_pytest.log
Q="--steps 8 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 8"
timeout 600 python bench.py $Q --handles 4 --lanes 592 > gpurun_out/r4_h4.log 2> gpurun_out/r4_h4.err || tail -5 gpurun_out/r4_h4.err
timeout 600 python bench.py $Q --handles 1 --lanes 592 --profile-json gpurun_out/r4_prof_h1_l592.json > gpurun_out/r4_h1.log 2> gpurun_out/r4_h1.err || tail -5 gpurun_out/r4_h1.err
python -c "
import json
for f in ('r4_h4','r4_h1'):
    l=json.loads(open('gpurun_out/%s.log'%f).read().strip().splitlines()[-1]); print(f, round(l['value']))
p=json.load(open('gpurun_out/r4_prof_h1_l592.json'))
k0=p['per_kernel_ms_over_profiled_step_handle0']
tot=sum(v['ms'] for v in k0.values()); print('sum', round(tot,1))
for k,v in sorted(k0.items(), key=lambda kv:-kv[1]['ms'])[:16]:
    print('%-28s %7.2f ms n=%4d avg %.3f ms'%(k,v['ms'],v['launches'],v['ms']/max(v['launches'],1)))
"
