# what the driver runs at the end of a round, in one gpurun call: the no-alloc test alone, the GPU suite, the default bench (+ event profile), the reference arm
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > gpurun_out/r1_gpu.txt; nproc >> gpurun_out/r1_gpu.txt; free -g | head -2 >> gpurun_out/r1_gpu.txt
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r1_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r1_pytest.log
timeout 1500 python bench.py --steps 20 --warmup 5 --profile-json gpurun_out/r1_prof.json > gpurun_out/r1_bench.log 2> gpurun_out/r1_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r1_bench.err
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r1_ref.log 2> gpurun_out/r1_ref.err; echo "ref rc=$?"
