# final build of round 2: ncu launch list of one steady-state stretch, then ncu --set full of the top kernels (296 lanes, one handle)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 4 --handles 1 --lanes 296"
$CMD > gpurun_out/fin_plain.log 2>&1 || { tail -3 gpurun_out/fin_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 5700 -c 1200 --csv --log-file gpurun_out/r02_final_ncu_launches_vlp16_l296.csv $CMD > gpurun_out/fin_list.log 2>&1
echo "list rc=$?"; wc -l gpurun_out/r02_final_ncu_launches_vlp16_l296.csv
K='regex:fa_odometry_kernel|fa_sort_sectors_kernel|fa_voxel_ring_kernel|ip_ccl_union_kernel|mo_lm_persistent_kernel|im_merge_old_kernel|im_reduce_kernel'
timeout 1500 ncu --set full --import-source on --clock-control none -k "$K" -s 620 -c 24 -f -o gpurun_out/r02_final_full_vlp16_l296 $CMD > gpurun_out/fin_full.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/fin_full.log
ncu -i gpurun_out/r02_final_full_vlp16_l296.ncu-rep --page raw --csv > gpurun_out/r02_final_full_vlp16_l296_raw.csv 2>/dev/null
ls -la gpurun_out/ | grep final
sz=$(stat -c %s gpurun_out/r02_final_full_vlp16_l296.ncu-rep); if [ "$sz" -gt 55000000 ]; then rm gpurun_out/r02_final_full_vlp16_l296.ncu-rep; echo "report too large, kept the csv only"; fi
