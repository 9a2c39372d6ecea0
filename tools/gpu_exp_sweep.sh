# lanes x handles sweep of the device-resident arm after the round-2 kernel changes
mkdir -p gpurun_out
Q="--steps 6 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 8"
for LH in "592 4" "592 6" "592 8" "888 6" "1184 8" "740 5" "444 3"; do
  set -- $LH
  timeout 600 python bench.py $Q --lanes $1 --handles $2 > gpurun_out/sw_$1_$2.log 2> gpurun_out/sw_$1_$2.err || { tail -3 gpurun_out/sw_$1_$2.err; continue; }
  python -c "
import json
l=json.loads(open('gpurun_out/sw_$1_$2.log').read().strip().splitlines()[-1])
print('lanes $1 handles $2 value', round(l['value']), 'ms/step', round(l['ms_per_step'],2))"
  nvidia-smi --query-gpu=memory.used --format=csv,noheader
done
