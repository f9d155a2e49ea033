# round-end evidence at HEAD: smoke, bench (both arms), ncu launch list and one full capture of the step's kernels
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r2v.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke_r2v.log; tail -n 2 gpurun_out/smoke_r2v.log
python bench.py --impl reference > gpurun_out/bench_ref_r2v.json 2> gpurun_out/bench_ref_r2v.err
python bench.py > gpurun_out/bench_r2v.json 2> gpurun_out/bench_r2v.err; echo "bench rc=$?"
python -c "import json;d=json.loads(open('gpurun_out/bench_r2v.json').read().strip().splitlines()[-1]);r=d['roofline'];e=d['e2e'];print(d['value'],d['ms_per_step'],r['frac'],r['step_frac'],e['value'],e['host_fill_gbs'],e['host_fill_ceiling_gbs'],d['also']['config1_single_sweep']['ms_per_call'],d['also']['config1_reference_symbol']['ms_per_call'])"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-also"
$CMD > gpurun_out/plain_r2v.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 16 -c 16 --csv --log-file gpurun_out/launches_r2v.csv $CMD > gpurun_out/ncu_l_r2v.log 2>&1
ncu --set full --clock-control none --import-source on -s 16 -c 4 -o gpurun_out/step_full_r2v $CMD > gpurun_out/ncu_f_r2v.log 2>&1
tail -n 2 gpurun_out/ncu_f_r2v.log
