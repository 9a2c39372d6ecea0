LEGO_NVCC_DEFS="-DLEGO_ODO_BRUTE_MAX=1000000" python -c "
import lego_loam_b200
from lego_loam_b200.build import build
build(force=True)" > gpurun_out/det_build.log 2>&1 || tail -5 gpurun_out/det_build.log
echo "brute-force nearest:"; python tools/fa_determinism.py 340 320 2>&1 | tail -4
