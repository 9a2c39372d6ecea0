mkdir -p gpurun_out
export LEGO_SHELL_BACKTRACE=1
timeout 900 python -m pytest tests/test_gpu_quirks.py tests/test_gpu_ros_shells.py tests/test_gpu_stage_api.py -x -q -s -m gpu > gpurun_out/dbg_shell_b.log 2>&1; echo "quirks+ros+stage rc=$?"; grep -n "shell_harness" -A12 gpurun_out/dbg_shell_b.log | head -20; tail -30 gpurun_out/dbg_shell_b.log | cut -c1-300
