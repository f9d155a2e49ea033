set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2i.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2i.log
tail -n 6 gpurun_out/pytest_gpu_r2i.log
: > gpurun_out/variants_r2i.txt
VARIANT_ENV="FUSED=default" timeout 300 bash tools/variant_bench.sh "" >> gpurun_out/variants_r2i.txt 2>&1
VARIANT_ENV="FUSED=0" MV3D_FUSED=0 timeout 300 bash tools/variant_bench.sh "" >> gpurun_out/variants_r2i.txt 2>&1
cat gpurun_out/variants_r2i.txt
timeout 600 python bench.py > gpurun_out/bench_r2i.json 2> gpurun_out/bench_r2i.err; echo "bench rc=$?"
head -c 1500 gpurun_out/bench_r2i.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-also"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:encode_fused -s 3 -c 1 -o gpurun_out/fused_full_r2i $CMD > gpurun_out/ncu_f_r2i.log 2>&1
tail -n 2 gpurun_out/ncu_f_r2i.log
