set -x
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
nproc > gpurun_out/r2f_n${N}_host.txt; lscpu | grep -E "Model name|Socket|NUMA|^CPU\(s\)" >> gpurun_out/r2f_n${N}_host.txt
$TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_r2f_n$N.json 2> gpurun_out/bench_r2f_n$N.err; tail -c 600 gpurun_out/bench_r2f_n$N.json; tail -3 gpurun_out/bench_r2f_n$N.err
MV3D_HOST_READBACK=dense $TR bench.py --gpus $N --steps 5 --warmup 3 --only-e2e > gpurun_out/e2e_dense_r2f_n$N.json 2>> gpurun_out/bench_r2f_n$N.err
MV3D_HOST_PIN=0 $TR bench.py --gpus $N --steps 5 --warmup 3 --only-e2e > gpurun_out/e2e_nopin_r2f_n$N.json 2>> gpurun_out/bench_r2f_n$N.err
MV3D_HOST_THREADS=8 $TR bench.py --gpus $N --steps 5 --warmup 3 --only-e2e > gpurun_out/e2e_t8_r2f_n$N.json 2>> gpurun_out/bench_r2f_n$N.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_r2f_n$N.json 2>> gpurun_out/bench_r2f_n$N.err
for f in gpurun_out/e2e_*_r2f_n$N.json; do python -c "
import json,sys; d=json.load(open('$f')); e=d['e2e']; print('$f', round(e['value']), round(e['host_fill_gbs'],1), round(e['host_fill_ceiling_gbs'],1), e['host'])"; done
