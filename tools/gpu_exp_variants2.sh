# candidates: 592x1 profile, 592x4 throughput, one-lane latency
mkdir -p gpurun_out
Q="--no-e2e --no-cpu-baseline --secondary none"
for V in default "$@"; do
  if [ "$V" = default ]; then unset LEGO_B200_LIB; else export LEGO_B200_LIB=$PWD/lego-loam_comment_b200/variants/liblego_b200_$V.so; fi
  timeout 600 python bench.py $Q --no-latency --steps 4 --warmup 5 --handles 1 --lanes 592 --profile-json gpurun_out/var_$V.json > gpurun_out/var_$V.log 2> gpurun_out/var_$V.err || tail -3 gpurun_out/var_$V.err
  timeout 600 python bench.py $Q --steps 8 --warmup 5 > gpurun_out/var4_$V.log 2> gpurun_out/var4_$V.err || tail -3 gpurun_out/var4_$V.err
  python -c "
import json
p=json.load(open('gpurun_out/var_$V.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
tot=sum(v['ms'] for v in k0.values()); print('%-10s 592x1 value %6d sum ms %7.2f'%('$V', round(p['line']['value']), tot), {k: round(v['ms'],2) for k,v in k0.items() if k in ('mo_lm_persistent_kernel','fa_odometry_kernel','fa_sort_sectors_kernel','fa_voxel_ring_kernel')})
l=json.loads(open('gpurun_out/var4_$V.log').read().strip().splitlines()[-1]); print('%-10s 592x4 value %6d'%('$V', round(l['value'])), 'latency', l['latency']['ms_per_scan'])"
done
