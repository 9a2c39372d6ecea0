# final build of round 2: ncu launch list of one steady-state stretch (296 lanes), then ncu --set full of the top kernels (148 lanes: the replay
# saves and restores the lanes' state, which is what a capture costs)
mkdir -p gpurun_out
Q="--steps 1 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 4 --handles 1"
python bench.py $Q --lanes 296 > gpurun_out/fin_plain.log 2>&1 || { tail -3 gpurun_out/fin_plain.log; exit 1; }
timeout 420 ncu --metrics gpu__time_duration.sum --clock-control none -s 5700 -c 800 --csv --log-file gpurun_out/r02_final_ncu_launches_vlp16_l296.csv python bench.py $Q --lanes 296 > gpurun_out/fin_list.log 2>&1
echo "list rc=$?"; wc -l gpurun_out/r02_final_ncu_launches_vlp16_l296.csv
K='regex:fa_odometry_kernel|fa_sort_sectors_kernel|fa_voxel_ring_kernel|ip_ccl_tile_kernel|mo_lm_persistent_kernel|im_merge_old_kernel|im_reduce_kernel'
timeout 420 ncu --set full --import-source on --clock-control none -k "$K" -s 620 -c 19 -f -o gpurun_out/r02_final_full_vlp16_l148 python bench.py $Q --lanes 148 > gpurun_out/fin_full.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/fin_full.log | cut -c1-200
ncu -i gpurun_out/r02_final_full_vlp16_l148.ncu-rep --page raw --csv > gpurun_out/r02_final_full_vlp16_l148_raw.csv 2>/dev/null
ls -la gpurun_out/ | grep final
sz=$(stat -c %s gpurun_out/r02_final_full_vlp16_l148.ncu-rep); if [ "$sz" -gt 40000000 ]; then rm gpurun_out/r02_final_full_vlp16_l148.ncu-rep; echo "report too large, kept the csv only"; fi
