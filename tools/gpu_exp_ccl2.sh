mkdir -p gpurun_out
Q="--steps 4 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none"
for TW in 256 128 64; do
LEGO_CCL_TW=$TW timeout 600 python bench.py $Q --handles 1 --lanes 592 --profile-json gpurun_out/ccltw$TW.json > gpurun_out/ccltw$TW.log 2> gpurun_out/ccltw$TW.err || tail -3 gpurun_out/ccltw$TW.err
python -c "
import json
p=json.load(open('gpurun_out/ccltw$TW.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
tot=sum(v['ms'] for v in k0.values()); print('TW=$TW value', round(p['line']['value']), 'sum ms', round(tot,2), {k: round(v['ms'],2) for k,v in k0.items() if k.startswith('ip_ccl')})"
done
