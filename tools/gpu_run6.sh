mkdir -p gpurun_out
Q="--steps 8 --warmup 5 --no-cpu-baseline --no-latency --secondary none --distinct 8"
for cfg in "0 4" "1 4" "1 8"; do
set -- $cfg
H=4; [ "$2" = "8" ] && H=8
timeout 600 python bench.py $Q --host-copies $1 --threads $2 --handles $H > gpurun_out/r6_hc$1_t$2.log 2> gpurun_out/r6_hc$1_t$2.err || tail -5 gpurun_out/r6_hc$1_t$2.err
python -c "
import json
l=json.loads(open('gpurun_out/r6_hc$1_t$2.log').read().strip().splitlines()[-1]); print('host_copies $1 threads $2 handles $H: device', round(l['value']), 'e2e', round(l['e2e']['value']), 'GB/s', l['e2e']['h2d_gbs_achieved'])"
done
