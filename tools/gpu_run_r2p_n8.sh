# 8-GPU box: both arms of the bench line at HEAD (run with gpurun --gpus 8)
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > gpurun_out/bench_ref_r2p_n8.json 2> gpurun_out/bench_ref_r2p_n8.err
timeout 900 $TR bench.py --gpus 8 > gpurun_out/bench_r2p_n8.json 2> gpurun_out/bench_r2p_n8.err; echo "bench rc=$?"
python -c "import json;d=json.loads(open('gpurun_out/bench_r2p_n8.json').read().strip().splitlines()[-1]);e=d['e2e'];print('N8 value',d['value'],'e2e',e['value'],e['host_fill_gbs'],e['host_fill_ceiling_gbs'],e['d2h_bytes_per_step']); print({k:(v.get('value') if isinstance(v,dict) else v) for k,v in d['also'].items()})"
MV3D_HOST_TRACE=1 timeout 300 $TR bench.py --gpus 8 --steps 3 --warmup 3 --only-e2e > gpurun_out/e2e_r2p_n8.json 2> gpurun_out/e2e_r2p_n8.err
grep "mv3d host" gpurun_out/e2e_r2p_n8.err | tail -n 8
tail -c 400 gpurun_out/bench_ref_r2p_n8.json
