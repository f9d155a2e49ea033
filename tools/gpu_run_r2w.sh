set -x
mkdir -p gpurun_out
MV3D_NVCC_EXTRA="-DMV3D_SELF_CHECK" python -c "from mv3d_b200 import _lib; _lib.build_library(force=True)"
timeout 900 python -m pytest tests/test_gpu_config4.py tests/test_gpu_errors.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_selfcheck_r2w.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_selfcheck_r2w.log
tail -n 4 gpurun_out/pytest_selfcheck_r2w.log
python -c "from mv3d_b200 import _lib; _lib.build_library(force=True)"
timeout 900 python -m pytest tests/test_gpu_config4.py tests/test_gpu_errors.py -m gpu -x -q > gpurun_out/pytest_r2w.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_r2w.log
tail -n 4 gpurun_out/pytest_r2w.log
