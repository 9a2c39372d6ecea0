# ncu launch list (device time of every launch, serialised, cold cache: compare SHARES) of one timed step at steady state
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 4 --handles 1 --lanes 296"
$CMD > gpurun_out/r02_list_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 7000 -c 1400 --csv --log-file gpurun_out/r02_ncu_launches_vlp16_l296.csv $CMD > gpurun_out/r02_list_ncu.log 2>&1
echo "rc=$?"; wc -l gpurun_out/r02_ncu_launches_vlp16_l296.csv; tail -2 gpurun_out/r02_list_ncu.log
