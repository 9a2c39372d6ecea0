# thread-block clusters for the odometry kernel: parity tests, then one-lane latency by cluster size and the saturated headline
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_feature_association.py tests/test_gpu_map_optimization.py tests/test_gpu_imu.py -x -q -m gpu 2>&1 | tail -5
Q="--steps 4 --warmup 5 --no-e2e --no-cpu-baseline --secondary none --distinct 1 --handles 1"
for CL in 1 2 4 8; do
  LEGO_ODO_CLUSTER=$CL timeout 600 python bench.py $Q --lanes 1 --profile-json gpurun_out/cl_$CL.json > gpurun_out/cl_$CL.log 2> gpurun_out/cl_$CL.err || tail -3 gpurun_out/cl_$CL.err
  python -c "
import json
p=json.load(open('gpurun_out/cl_$CL.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
print('CL=$CL lanes 1 latency', {k: round(v,3) for k,v in p['line']['latency']['ms_per_scan'].items()}, 'odometry ms/25 scans', round(k0['fa_odometry_kernel']['ms'],3))"
done
for LN in 16 64; do for CL in 1 0; do
  LEGO_ODO_CLUSTER=$CL timeout 600 python bench.py $Q --no-latency --distinct 8 --lanes $LN --profile-json gpurun_out/cl_l${LN}_$CL.json > gpurun_out/cl_l${LN}_$CL.log 2> gpurun_out/cl_l${LN}_$CL.err || tail -3 gpurun_out/cl_l${LN}_$CL.err
  python -c "
import json
p=json.load(open('gpurun_out/cl_l${LN}_$CL.json')); k0=p['per_kernel_ms_over_profiled_step_handle0']
print('CL=$CL (0 = auto) lanes $LN value', round(p['line']['value']), 'odometry ms/25 scans', round(k0['fa_odometry_kernel']['ms'],3))"
done; done
