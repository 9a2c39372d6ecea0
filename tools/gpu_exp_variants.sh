# compile-time variants built locally into lego-loam_comment_b200/variants/ (LEGO_B200_LIB selects one): per-kernel profile, 592 lanes x 1 handle
mkdir -p gpurun_out
Q="--steps 4 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none"
for V in default "$@"; do
  if [ "$V" = default ]; then unset LEGO_B200_LIB; else export LEGO_B200_LIB=$PWD/lego-loam_comment_b200/variants/liblego_b200_$V.so; fi
  timeout 600 python bench.py $Q --handles 1 --lanes 592 --profile-json gpurun_out/var_$V.json > gpurun_out/var_$V.log 2> gpurun_out/var_$V.err || tail -3 gpurun_out/var_$V.err
  python -c "
import json
p=json.load(open('gpurun_out/var_$V.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
tot=sum(v['ms'] for v in k0.values()); print('%-10s value %6d sum ms %7.2f'%('$V', round(p['line']['value']), tot), {k: round(v['ms'],2) for k,v in k0.items() if k in ('mo_lm_persistent_kernel','fa_odometry_kernel','fa_voxel_ring_kernel','im_merge_old_kernel','ip_ccl_tile_kernel','im_sort_arrivals','vg_filter(currentScan)','vg_filter(surfTotal)')})"
done
