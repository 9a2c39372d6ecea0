# usage: bash tools/gpu_sanitize.sh memcheck|racecheck   (ONE tool per gpurun call, B200_PROFILING.md)
mkdir -p gpurun_out
TOOL=$1
python tools/sanitize_target.py 8 > gpurun_out/r02_sanitize_plain.log 2>&1 &&
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 30 --log-file gpurun_out/r02_sanitizer_$TOOL.txt python tools/sanitize_target.py 8 > gpurun_out/r02_sanitize_$TOOL.out 2>&1
echo "rc=$?"; tail -3 gpurun_out/r02_sanitize_$TOOL.out; tail -8 gpurun_out/r02_sanitizer_$TOOL.txt
