# CUDA-graph replay of the per-scan launch sequences: parity A/B test, one-lane latency and saturated throughput with and without
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pipeline_api.py -x -q -m gpu 2>&1 | tail -5
QL="--steps 4 --warmup 5 --no-e2e --no-cpu-baseline --secondary none --distinct 1 --lanes 1 --handles 1"
for G in 0 1; do
  LEGO_GRAPH=$G timeout 600 python bench.py $QL > gpurun_out/graph_lat_g$G.log 2> gpurun_out/graph_lat_g$G.err || tail -3 gpurun_out/graph_lat_g$G.err
  python -c "
import json
l=json.loads(open('gpurun_out/graph_lat_g$G.log').read().strip().splitlines()[-1])
print('G=$G one lane value', round(l['value']), 'latency', json.dumps(l.get('latency')))"
done
QT="--steps 8 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none"
for G in 0 1; do
  LEGO_GRAPH=$G timeout 600 python bench.py $QT > gpurun_out/graph_tp_g$G.log 2> gpurun_out/graph_tp_g$G.err || tail -3 gpurun_out/graph_tp_g$G.err
  python -c "
import json
l=json.loads(open('gpurun_out/graph_tp_g$G.log').read().strip().splitlines()[-1])
print('G=$G 592x4 value', round(l['value']), 'ms/step', round(l['ms_per_step'],2), 'host submit', l.get('host_submit_ms_per_step'))"
done
for L in 16 64; do for G in 0 1; do
  LEGO_GRAPH=$G timeout 600 python bench.py $QT --lanes $L --handles 1 --distinct 8 > gpurun_out/graph_l${L}_g$G.log 2> gpurun_out/graph_l${L}_g$G.err || tail -3 gpurun_out/graph_l${L}_g$G.err
  python -c "
import json
l=json.loads(open('gpurun_out/graph_l${L}_g$G.log').read().strip().splitlines()[-1])
print('G=$G lanes $L x1 value', round(l['value']), 'ms/step', round(l['ms_per_step'],2))"
done; done
