set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2q.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2q.log
tail -n 6 gpurun_out/pytest_gpu_r2q.log
nproc
MV3D_HOST_TRACE=1 timeout 300 python bench.py --steps 5 --warmup 3 --only-e2e > gpurun_out/e2e_r2q.json 2> gpurun_out/e2e_r2q.err
grep "mv3d host" gpurun_out/e2e_r2q.err | tail -n 3
python -c "import json;d=json.loads(open('gpurun_out/e2e_r2q.json').read().strip().splitlines()[-1]);e=d['e2e'];print(e['value'],e['host_fill_gbs'],e['host_fill_ceiling_gbs'])"
timeout 600 python bench.py > gpurun_out/bench_r2q.json 2> gpurun_out/bench_r2q.err; echo "bench rc=$?"
python -c "import json;d=json.loads(open('gpurun_out/bench_r2q.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['e2e']['value'],d['roofline']['frac'],d['roofline']['step_frac'])"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_r2q.json 2> gpurun_out/bench_ref_r2q.err; tail -c 600 gpurun_out/bench_ref_r2q.json
