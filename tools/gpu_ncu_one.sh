# ncu --set full of ONE kernel (regex $1, skip $2 launches, capture $3) on a short 592-lane run; raw + source pages come back as csv
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 4 --handles 1 --lanes 592"
timeout 900 ncu --set full --import-source on --clock-control none -k "regex:$1" -s ${2:-10} -c ${3:-2} -f -o gpurun_out/one $CMD > gpurun_out/one.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/one.log | cut -c1-300
ncu -i gpurun_out/one.ncu-rep --page raw --csv > gpurun_out/one_raw.csv 2>/dev/null
ncu -i gpurun_out/one.ncu-rep --page source --csv > gpurun_out/one_source.csv 2>/dev/null
rm -f gpurun_out/one.ncu-rep
ls -la gpurun_out/one*
