# bench line at N GPUs of one box: bash tools/gpu_run_r2x.sh N   (run with gpurun --gpus N)
set -x
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
timeout 600 $TR bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/bench_ref_r2x_n$N.json 2> gpurun_out/bench_ref_r2x_n$N.err
timeout 900 $TR bench.py --gpus $N > gpurun_out/bench_r2x_n$N.json 2> gpurun_out/bench_r2x_n$N.err; echo "bench rc=$?"
python -c "import json;d=json.loads(open('gpurun_out/bench_r2x_n$N.json').read().strip().splitlines()[-1]);e=d['e2e'];print('N=$N value',d['value'],d['ms_per_step'],'e2e',e['value'],e['host_fill_gbs'],e['host_fill_ceiling_gbs'],e['host']); print({k:(v.get('value') if isinstance(v,dict) else v) for k,v in d['also'].items()})"
python -c "import json;d=json.loads(open('gpurun_out/bench_ref_r2x_n$N.json').read().strip().splitlines()[-1]);print('ref',d['value'],d['cpu_baseline']['cores'])"
