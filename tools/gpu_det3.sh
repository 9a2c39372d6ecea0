for T in 32 64; do
LEGO_NVCC_DEFS="-DLEGO_ODO_THREADS=$T" python -c "
import lego_loam_b200
from lego_loam_b200.build import build
build(force=True)" > gpurun_out/det_build.log 2>&1 || tail -5 gpurun_out/det_build.log
echo "ODO_THREADS=$T:"; python tools/fa_determinism.py 332 324 2>&1 | grep -v "pipeline\|transformCur \[" | tail -3
done
