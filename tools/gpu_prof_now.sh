# per-kernel CUDA-event profile of the current build: 592 lanes on one handle (steady-state window), and one lane alone
mkdir -p gpurun_out
Q="--steps 4 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none"
timeout 600 python bench.py $Q --handles 1 --lanes 592 --profile-json gpurun_out/now_l592h1.json > gpurun_out/now_l592h1.log 2> gpurun_out/now_l592h1.err || tail -3 gpurun_out/now_l592h1.err
timeout 600 python bench.py $Q --handles 1 --lanes 1 --distinct 1 --profile-json gpurun_out/now_l1.json > gpurun_out/now_l1.log 2> gpurun_out/now_l1.err || tail -3 gpurun_out/now_l1.err
for f in now_l592h1 now_l1; do python -c "
import json
p=json.load(open('gpurun_out/$f.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
tot=sum(v['ms'] for v in k0.values()); print('$f value', round(p['line']['value']), 'sum ms', round(tot,2))
for k,v in sorted(k0.items(), key=lambda kv:-kv[1]['ms'])[:24]: print('   %-32s %7.3f ms %5.1f%% n=%s'%(k,v['ms'],100*v['ms']/tot,v.get('launches')))"
done
