# one-lane latency and its per-kernel composition, launch loop vs persistent LM kernel
mkdir -p gpurun_out
Q="--steps 4 --warmup 5 --no-e2e --no-cpu-baseline --secondary none --distinct 1 --lanes 1 --handles 1"
for P in 0 1; do
  LEGO_LM_PERSISTENT=$P timeout 600 python bench.py $Q --profile-json gpurun_out/lat_p$P.json > gpurun_out/lat_p$P.log 2> gpurun_out/lat_p$P.err || tail -3 gpurun_out/lat_p$P.err
  python -c "
import json
p=json.load(open('gpurun_out/lat_p$P.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
print('P=$P latency', json.dumps(p['line'].get('latency')))
tot=sum(v['ms'] for v in k0.values()); print(' sum ms over 25 scans', round(tot,2))
for k,v in sorted(k0.items(), key=lambda kv:-kv[1]['ms'])[:16]: print('   %-32s %6.3f ms n=%s'%(k,v['ms'],v.get('launches')))"
done
