# BASELINE.json config 5: 8 independent 2000-scan sequences with loop closure, on N GPUs of one box (N = $1), 8/N sequences per GPU.
# Writes profiles-ready JSON lines to gpurun_out/r02_config5_n$N*.json
N=${1:-1}
mkdir -p gpurun_out
LANES=$((8 / N))
ARGS="--workload loop --lanes $LANES --handles 1 --distinct $LANES --steps 75 --warmup 5 --secondary none --no-latency --gpus $N"
if [ "$N" = "1" ]; then
  timeout 700 python bench.py $ARGS > gpurun_out/r02_config5_n1.json 2> gpurun_out/r02_config5_n1.err || tail -c 800 gpurun_out/r02_config5_n1.err
  timeout 700 python bench.py $ARGS --impl reference > gpurun_out/r02_config5_reference.json 2> gpurun_out/r02_config5_reference.err || tail -c 800 gpurun_out/r02_config5_reference.err
  # the same workload with the GPU filled, first 400 scans only (4 handles x 64 sequences, 16 ICP slots per handle): beyond a
  # few hundred key frames with loop factors the host-side pose-graph solve of 256 sequences dominates everything
  timeout 600 python bench.py --workload loop --lanes 256 --handles 4 --distinct 8 --steps 11 --warmup 5 --secondary none --no-latency --no-cpu-baseline > gpurun_out/r02_config5_saturated400_n1.json 2> gpurun_out/r02_config5_saturated400_n1.err || tail -c 800 gpurun_out/r02_config5_saturated400_n1.err
else
  timeout 700 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py $ARGS --no-cpu-baseline > gpurun_out/r02_config5_n$N.json 2> gpurun_out/r02_config5_n$N.err || tail -c 800 gpurun_out/r02_config5_n$N.err
fi
python - <<PY
import json, glob
for f in sorted(glob.glob('gpurun_out/r02_config5*n$N*.json') + (glob.glob('gpurun_out/r02_config5_reference.json') if '$N' == '1' else [])):
    try:
        l = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, 'value %.0f' % l['value'], 'e2e %.0f' % l['e2e']['value'], l['config'].get('timed_scan_range'), 'lanes', l['config'].get('lanes_per_gpu'), 'n_gpus', l['n_gpus'],
              'cpu1 %.1f' % l['cpu_baseline']['value'] if 'cpu_baseline' in l else '')
    except Exception as e:
        print(f, 'unreadable', e)
PY
