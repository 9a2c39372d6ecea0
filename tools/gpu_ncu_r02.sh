# ncu --set full of the top kernels at the steady-state local map (scans 128..131 of every lane; mapping call at scan 130)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 5 --no-e2e --no-cpu-baseline --no-latency --secondary none --distinct 4 --handles 1 --lanes 296"
K='regex:fa_odometry_kernel|fa_sort_sectors_kernel|fa_voxel_ring_kernel|ip_ccl_union_kernel|mo_search_kernel|mo_residual_kernel|im_merge_old_kernel|im_reduce_kernel'
$CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none -k "$K" -s 1216 -c 40 -f -o gpurun_out/r02_full_top8_vlp16_l296 $CMD > gpurun_out/r02_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r02_ncu.log; ls -la gpurun_out/*.ncu-rep
