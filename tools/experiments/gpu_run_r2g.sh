set -x
mkdir -p gpurun_out
export CUDA_VISIBLE_DEVICES_ORIG=$CUDA_VISIBLE_DEVICES
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_config4.py tests/test_gpu_pytwin.py tests/test_gpu_fuzz.py tests/test_gpu_clidar.py -m gpu -x -q > gpurun_out/pytest_fused_r2g.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_fused_r2g.log
tail -8 gpurun_out/pytest_fused_r2g.log
: > gpurun_out/variants_r2g.txt
VARIANT_ENV="FUSED=1" timeout 300 bash tools/variant_bench.sh "" >> gpurun_out/variants_r2g.txt 2>&1
VARIANT_ENV="FUSED=0" MV3D_FUSED=0 timeout 300 bash tools/variant_bench.sh "" >> gpurun_out/variants_r2g.txt 2>&1
cat gpurun_out/variants_r2g.txt
# host trace of the e2e leg on the multi-GPU box's host
MV3D_HOST_TRACE=1 python bench.py --steps 3 --warmup 3 --only-e2e > gpurun_out/e2e_trace_n1_r2g.json 2> gpurun_out/e2e_trace_n1_r2g.err
MV3D_HOST_TRACE=1 MV3D_HOST_THREADS=4 python bench.py --steps 3 --warmup 3 --only-e2e > gpurun_out/e2e_trace_n1_t4_r2g.json 2> gpurun_out/e2e_trace_n1_t4_r2g.err
MV3D_HOST_TRACE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 --only-e2e > gpurun_out/e2e_trace_n2_r2g.json 2> gpurun_out/e2e_trace_n2_r2g.err
grep "mv3d host" gpurun_out/e2e_trace_n1_r2g.err | tail -2; grep "mv3d host" gpurun_out/e2e_trace_n1_t4_r2g.err | tail -2; grep "mv3d host" gpurun_out/e2e_trace_n2_r2g.err | tail -3
nproc
