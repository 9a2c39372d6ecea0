mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pipeline_api.py -x -q -m gpu 2>&1 | tail -2
Q="--steps 8 --warmup 5 --no-cpu-baseline --no-latency --secondary none --distinct 8"
for hc in 0 2; do
timeout 600 python bench.py $Q --host-copies $hc > gpurun_out/r7_hc$hc.log 2> gpurun_out/r7_hc$hc.err || tail -5 gpurun_out/r7_hc$hc.err
python -c "
import json
l=json.loads(open('gpurun_out/r7_hc$hc.log').read().strip().splitlines()[-1]); print('host_copies $hc: device', round(l['value']), 'e2e', round(l['e2e']['value']), 'GB/s', l['e2e']['h2d_gbs_achieved'], 'plain', l['e2e']['pcie_h2d_gbs_plain_copy'])"
done
