# host diagnostics of the e2e leg on the multi-GPU box (run with gpurun --gpus 2)
set -x
mkdir -p gpurun_out
O=gpurun_out/host_r2j.txt
{ nproc; lscpu | grep -i "model name\|socket\|numa\|thread\|core\|L2\|L3\|flags" | cut -c1-400; for c in 0 1 2 3 16 17; do echo -n "cpu$c siblings: "; cat /sys/devices/system/cpu/cpu$c/topology/thread_siblings_list; done; free -g | head -2; cat /sys/kernel/mm/transparent_hugepage/enabled; nvidia-smi topo -m; } > $O 2>&1
nvcc -O3 -std=c++17 -o /tmp/host_bw tools/host_bw.cu -lpthread 2>/dev/null && /tmp/host_bw >> $O 2>&1
python tools/host_rebuild_list.py /tmp/list.bin >> $O 2>&1
nvcc -O3 -std=c++17 -o /tmp/host_rebuild_bench tools/host_rebuild_bench.cu mv3d_b200/csrc/host_expand.cu -lpthread 2>/dev/null && /tmp/host_rebuild_bench /tmp/list.bin 16 >> $O 2>&1
# two rebuild benches at once, 16 threads each: does the box take both?
( taskset -c 0-15 /tmp/host_rebuild_bench /tmp/list.bin 16 > /tmp/rb_a.txt 2>&1 & taskset -c 16-31 /tmp/host_rebuild_bench /tmp/list.bin 16 > /tmp/rb_b.txt 2>&1 & wait )
{ echo "--- two at once (cpus 0-15 | 16-31)"; cat /tmp/rb_a.txt; echo ---; cat /tmp/rb_b.txt; } >> $O
run() { # name, env..., nproc
  name=$1; np=$2; shift 2
  if [ "$np" = 1 ]; then env MV3D_HOST_TRACE=1 "$@" python bench.py --steps 3 --warmup 3 --only-e2e > gpurun_out/e2e_${name}_r2j.json 2> gpurun_out/e2e_${name}_r2j.err
  else env MV3D_HOST_TRACE=1 "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $np --steps 3 --warmup 3 --only-e2e > gpurun_out/e2e_${name}_r2j.json 2> gpurun_out/e2e_${name}_r2j.err; fi
  echo "== $name" >> $O; grep "mv3d host" gpurun_out/e2e_${name}_r2j.err | tail -n 2 >> $O
  python -c "import json;d=json.loads(open('gpurun_out/e2e_${name}_r2j.json').read().strip().splitlines()[-1]);e=d['e2e'];print(e['value'],e['host_fill_gbs'],e['host_fill_ceiling_gbs'],e['host'])" >> $O 2>&1
}
run n1_t16 1 A=1
run n1_t32 1 MV3D_HOST_THREADS=32
run n1_t8 1 MV3D_HOST_THREADS=8
run n2_t16 2 A=1
run n2_t8 2 MV3D_HOST_THREADS=8
run n2_t16_nopin 2 MV3D_HOST_PIN=0
run n2_dense 2 MV3D_HOST_READBACK=dense
cat $O
