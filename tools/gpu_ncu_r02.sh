# ncu --set full of the top kernels at the steady-state local map (scans 128..130 of every lane; mapping call at scan 129 or 130)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 4 --handles 1 --lanes 296"
K='regex:fa_odometry_kernel|fa_sort_sectors_kernel|fa_voxel_ring_kernel|ip_ccl_union_kernel|mo_search_kernel|im_merge_old_kernel|im_reduce_kernel'
$CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none -k "$K" -s 896 -c 24 -f -o gpurun_out/r02_full_top7_vlp16_l296 $CMD > gpurun_out/r02_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r02_ncu.log
ncu -i gpurun_out/r02_full_top7_vlp16_l296.ncu-rep --page raw --csv > gpurun_out/r02_full_top7_vlp16_l296_raw.csv 2>/dev/null
ls -la gpurun_out/
sz=$(stat -c %s gpurun_out/r02_full_top7_vlp16_l296.ncu-rep); if [ "$sz" -gt 55000000 ]; then rm gpurun_out/r02_full_top7_vlp16_l296.ncu-rep; echo "report too large, kept the csv only"; fi
