set -x
mkdir -p gpurun_out
nvcc -O3 -std=c++17 -o /tmp/host_rebuild_bench tools/host_rebuild_bench.cu mv3d_b200/csrc/host_expand.cu -lpthread 2>/dev/null; python tools/host_rebuild_list.py /tmp/list.bin
{ HUGE=0 /tmp/host_rebuild_bench /tmp/list.bin 16; HUGE=1 /tmp/host_rebuild_bench /tmp/list.bin 16; } > gpurun_out/host_rebuild_r2y.txt 2>&1
cat gpurun_out/host_rebuild_r2y.txt
MV3D_NVCC_EXTRA="-DMV3D_SELF_CHECK" python -c "from mv3d_b200 import _lib; _lib.build_library(force=True)"
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "many_planes or host_batch" > gpurun_out/pytest_planes_r2y.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_planes_r2y.log
tail -n 12 gpurun_out/pytest_planes_r2y.log
python -c "from mv3d_b200 import _lib; _lib.build_library(force=True)"
