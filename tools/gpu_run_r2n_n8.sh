# 8-GPU box: full bench line at N=8, then the host side of the e2e leg taken apart (run with gpurun --gpus 8)
set -x
mkdir -p gpurun_out
O=gpurun_out/host_r2n_n8.txt
{ nproc; lscpu | grep -i "model name\|socket\|numa\|thread\|core(s)\|L2\|L3"; grep -o 'avx512f\|avx2' /proc/cpuinfo | sort | uniq -c; free -g | head -2; } > $O 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
timeout 900 $TR bench.py --gpus 8 > gpurun_out/bench_r2n_n8.json 2> gpurun_out/bench_r2n_n8.err; echo "bench rc=$?" >> $O
python -c "import json;d=json.loads(open('gpurun_out/bench_r2n_n8.json').read().strip().splitlines()[-1]);print('N8 value',d['value'],'e2e',d['e2e']['value'],d['e2e']['host_fill_gbs'],d['e2e']['host_fill_ceiling_gbs']); print({k:(v.get('value') if isinstance(v,dict) else v) for k,v in d['also'].items()})" >> $O 2>&1
run() { name=$1; shift
  env MV3D_HOST_TRACE=1 "$@" timeout 300 $TR bench.py --gpus 8 --steps 3 --warmup 3 --only-e2e > gpurun_out/e2e_${name}_r2n.json 2> gpurun_out/e2e_${name}_r2n.err
  echo "== $name" >> $O; grep "mv3d host" gpurun_out/e2e_${name}_r2n.err | tail -n 8 >> $O
  python -c "import json;d=json.loads(open('gpurun_out/e2e_${name}_r2n.json').read().strip().splitlines()[-1]);e=d['e2e'];print(e['value'],e['host_fill_gbs'],e['host_fill_ceiling_gbs'],e['host']['fill_threads_per_rank'],e['host']['pinned_to_cpu_share'])" >> $O 2>&1
}
run n8_default A=1
run n8_t3 MV3D_HOST_THREADS=3
run n8_t2 MV3D_HOST_THREADS=2
run n8_t6_nopin MV3D_HOST_THREADS=6 MV3D_HOST_PIN=0
# the rebuild alone, no CUDA: 8 processes x 4 pinned CPUs at once, then one process on all CPUs
python tools/host_rebuild_list.py /tmp/list.bin >> $O 2>&1
nvcc -O3 -std=c++17 -o /tmp/host_rebuild_bench tools/host_rebuild_bench.cu mv3d_b200/csrc/host_expand.cu -lpthread 2>/dev/null
for r in 0 1 2 3 4 5 6 7; do taskset -c $((4*r))-$((4*r+3)) /tmp/host_rebuild_bench /tmp/list.bin 16 > /tmp/rb_$r.txt 2>&1 & done; wait
{ echo "--- rebuild bench, 8 processes x 4 pinned CPUs at once"; for r in 0 3 7; do echo "proc $r"; cat /tmp/rb_$r.txt; done; echo "--- one process, all CPUs"; /tmp/host_rebuild_bench /tmp/list.bin 16; } >> $O 2>&1
# one GPU on this box
MV3D_HOST_TRACE=1 timeout 300 python bench.py --steps 3 --warmup 3 --only-e2e > gpurun_out/e2e_n1_r2n.json 2> gpurun_out/e2e_n1_r2n.err
echo "== n1 on the 8-GPU box" >> $O; grep "mv3d host" gpurun_out/e2e_n1_r2n.err | tail -n 2 >> $O
python -c "import json;d=json.loads(open('gpurun_out/e2e_n1_r2n.json').read().strip().splitlines()[-1]);e=d['e2e'];print(e['value'],e['host_fill_gbs'],e['host_fill_ceiling_gbs'])" >> $O 2>&1
cat $O
