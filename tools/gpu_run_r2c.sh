set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2c.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2c.log
tail -5 gpurun_out/pytest_gpu_r2c.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r2c.json 2> gpurun_out/bench_r2c.err; tail -c 2500 gpurun_out/bench_r2c.json; tail -5 gpurun_out/bench_r2c.err
: > gpurun_out/variants_r2c.txt
for w in 0 1 2 4; do VARIANT_ENV="WAVES=$w" MV3D_POINT_WAVES=$w bash tools/variant_bench.sh "" >> gpurun_out/variants_r2c.txt 2>&1; done
VARIANT_ENV="PDL=0" MV3D_PDL=0 bash tools/variant_bench.sh "" >> gpurun_out/variants_r2c.txt 2>&1
VARIANT_ENV="PDL=1" bash tools/variant_bench.sh "" "-DMV3D_CLASSIFY_MIN_BLOCKS=6" "-DMV3D_POINT_ITEMS=4" "-DMV3D_POINT_ITEMS=1" "" >> gpurun_out/variants_r2c.txt 2>&1
cat gpurun_out/variants_r2c.txt
python tools/strategy_profile.py > gpurun_out/strategy_plain_r2c.log 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/strategies_full_r2c python tools/strategy_profile.py > gpurun_out/ncu_s_r2c.log 2>&1
cat gpurun_out/strategy_plain_r2c.log; tail -3 gpurun_out/ncu_s_r2c.log
