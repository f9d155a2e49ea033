set -x
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone_r2e.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_noalloc_alone_r2e.log
tail -4 gpurun_out/pytest_noalloc_alone_r2e.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2e.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2e.log
tail -5 gpurun_out/pytest_gpu_r2e.log
