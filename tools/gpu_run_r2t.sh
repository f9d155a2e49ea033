# the GPU suite on the -DMV3D_SELF_CHECK build (the library's own bounds checks; compute-sanitizer is closed on this pool)
set -x
mkdir -p gpurun_out
MV3D_NVCC_EXTRA="-DMV3D_SELF_CHECK" python -c "from mv3d_b200 import _lib; _lib.build_library(force=True)"
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_selfcheck_r2t.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_selfcheck_r2t.log
tail -n 6 gpurun_out/pytest_gpu_selfcheck_r2t.log
python - <<'PY' >> gpurun_out/pytest_gpu_selfcheck_r2t.log 2>&1
import ctypes
from mv3d_b200 import _lib
lib = _lib.load_library(); f = ctypes.c_uint(0)
print("self-check compiled in:", lib.mv3d_debug_selfcheck(None, ctypes.byref(f)) != 0)
PY
tail -n 1 gpurun_out/pytest_gpu_selfcheck_r2t.log
python -c "from mv3d_b200 import _lib; _lib.build_library(force=True)"
