set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2d.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2d.log
tail -5 gpurun_out/pytest_gpu_r2d.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r2d.json 2> gpurun_out/bench_r2d.err; tail -c 1500 gpurun_out/bench_r2d.json; tail -5 gpurun_out/bench_r2d.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_r2d.json 2>> gpurun_out/bench_r2d.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-also"
$CMD > gpurun_out/plain_r2d.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 16 -c 16 --csv --log-file gpurun_out/launches_r2d.csv $CMD > gpurun_out/ncu_l_r2d.log 2>&1
python tools/strategy_profile.py > gpurun_out/strategy_plain_r2d.log 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/strategies_full_r2d python tools/strategy_profile.py > gpurun_out/ncu_s_r2d.log 2>&1
cat gpurun_out/strategy_plain_r2d.log; tail -3 gpurun_out/ncu_s_r2d.log
