# 16-CTA odometry clusters: cluster-size equality test + pipeline API tests, then one-lane latency with clusters of 8 and 16
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_feature_association.py tests/test_gpu_pipeline_api.py -x -q -m gpu 2>&1 | tail -4
for CL in 8 16; do
LEGO_ODO_CLUSTER=$CL timeout 600 python bench.py --steps 2 --warmup 5 --no-e2e --no-cpu-baseline --secondary none --distinct 1 --lanes 1 --handles 1 > gpurun_out/cl$CL.log 2> gpurun_out/cl$CL.err || tail -3 gpurun_out/cl$CL.err
python -c "
import json
l=json.loads(open('gpurun_out/cl$CL.log').read().strip().splitlines()[-1]); print('cluster $CL', l['latency']['ms_per_scan'])"
done
