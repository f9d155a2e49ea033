set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2u.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2u.log
tail -n 4 gpurun_out/pytest_gpu_r2u.log
: > gpurun_out/configs_r2u.txt
for items in 2 4 3; do
  MV3D_NVCC_EXTRA="-DMV3D_ATOMIC_ITEMS=$items" python -c "from mv3d_b200 import _lib; _lib.build_library(force=True)"
  echo "== MV3D_ATOMIC_ITEMS=$items" >> gpurun_out/configs_r2u.txt
  timeout 300 python tools/configs_bench.py --steps 20 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print({k:(round(v,4) if isinstance(v,float) else v) for k,v in d.items() if k in ('case','config','strategy','ms_per_call','ms','value','sweeps_per_s','clouds_per_s','kernels_ms','kernel_ms')})" >> gpurun_out/configs_r2u.txt
done
MV3D_PDL=0 timeout 300 python tools/configs_bench.py --steps 20 2>/dev/null | head -c 3000 > gpurun_out/configs_nopdl_r2u.txt
python -c "from mv3d_b200 import _lib; _lib.build_library(force=True)"
cat gpurun_out/configs_r2u.txt | cut -c1-400
