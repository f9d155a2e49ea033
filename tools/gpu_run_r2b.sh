set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2b.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2b.log
tail -5 gpurun_out/pytest_gpu_r2b.log
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r2b.json 2> gpurun_out/bench_r2b.err; tail -c 1500 gpurun_out/bench_r2b.json
VARIANT_ENV="PDL=1" bash tools/variant_bench.sh "" > gpurun_out/variants_r2b.txt 2>&1
VARIANT_ENV="PDL=0" MV3D_PDL=0 bash tools/variant_bench.sh "" >> gpurun_out/variants_r2b.txt 2>&1
VARIANT_ENV="PDL=1" bash tools/variant_bench.sh "" >> gpurun_out/variants_r2b.txt 2>&1
cat gpurun_out/variants_r2b.txt
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-also"
$CMD > gpurun_out/plain_r2b.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 16 -c 12 --csv --log-file gpurun_out/launches_r2b.csv $CMD > gpurun_out/ncu_l_r2b.log 2>&1
$CMD > gpurun_out/plain_r2b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'classify|bin_kernel|tile_scan' -s 12 -c 3 -o gpurun_out/points_full_r2b $CMD > gpurun_out/ncu_p_r2b.log 2>&1
tail -3 gpurun_out/ncu_p_r2b.log
