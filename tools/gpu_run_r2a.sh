set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2a_gpu.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2a.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2a.log
tail -5 gpurun_out/pytest_gpu_r2a.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r2a.json 2> gpurun_out/bench_r2a.err; tail -c 3000 gpurun_out/bench_r2a.json
bash tools/variant_bench.sh "" "-DMV3D_CLASSIFY_MIN_BLOCKS=6" "-DMV3D_CLASSIFY_MIN_BLOCKS=5" "-DMV3D_POINT_ITEMS=4" "-DMV3D_POINT_ITEMS=1" "" > gpurun_out/variants_r2a.txt 2>&1
for tb in 30720 36864 40960 46080; do VARIANT_ENV="TILE_BYTES=$tb" MV3D_TILE_BYTES=$tb bash tools/variant_bench.sh "" >> gpurun_out/variants_r2a.txt 2>&1; done
cat gpurun_out/variants_r2a.txt
