mkdir -p gpurun_out
Q="--workload loop --distinct 8 --steps 11 --warmup 5 --no-latency --secondary none --no-cpu-baseline"
for cfg in "4 256 1" "4 256 4" "1 64 1"; do
  set -- $cfg
  timeout 900 python bench.py $Q --handles $1 --lanes $2 --threads $3 > gpurun_out/loop_h$1_l$2_t$3.log 2> gpurun_out/loop_h$1_l$2_t$3.err || tail -c 600 gpurun_out/loop_h$1_l$2_t$3.err
  python -c "
import json
l=json.loads(open('gpurun_out/loop_h$1_l$2_t$3.log').read().strip().splitlines()[-1])
print('handles $1 lanes $2 threads $3:', round(l['value']), 'e2e', round(l['e2e']['value']), 'dominant', l['roofline']['kernel'], round(l['roofline']['share_of_step'],3))"
done
