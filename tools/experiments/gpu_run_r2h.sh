set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_config4.py tests/test_gpu_pytwin.py tests/test_gpu_fuzz.py -m gpu -x -q > gpurun_out/pytest_fused_r2h.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_fused_r2h.log
tail -4 gpurun_out/pytest_fused_r2h.log
: > gpurun_out/variants_r2h.txt
VARIANT_ENV="FUSED=1" timeout 300 bash tools/variant_bench.sh "" >> gpurun_out/variants_r2h.txt 2>&1
VARIANT_ENV="FUSED=0" MV3D_FUSED=0 timeout 300 bash tools/variant_bench.sh "" >> gpurun_out/variants_r2h.txt 2>&1
VARIANT_ENV="FUSED=1" timeout 300 bash tools/variant_bench.sh "-DMV3D_FUSED_CLASSIFY_CHUNK=2048 -DMV3D_FUSED_BIN_CHUNK=4096" "-DMV3D_FUSED_CLASSIFY_CHUNK=512 -DMV3D_FUSED_BIN_CHUNK=1024" "" >> gpurun_out/variants_r2h.txt 2>&1
cat gpurun_out/variants_r2h.txt
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-also"
$CMD > gpurun_out/plain_r2h.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:encode_fused -s 3 -c 1 -o gpurun_out/fused_full_r2h $CMD > gpurun_out/ncu_f_r2h.log 2>&1
tail -2 gpurun_out/ncu_f_r2h.log
