#!/bin/bash
# Sweep the PATCH strategy in its copy-engine form (cudaMemsetAsync zero-fill on the side stream + tile_patch_kernel):
# arguments are <tile bytes>:<sweeps per fill chunk>. Run on the GPU box:  bash tools/ce_patch_sweep.sh 36864:8 18432:16
for cfg in "$@"; do
  tb=${cfg%:*}; ch=${cfg#*:}
  MV3D_TILE_BYTES=$tb MV3D_FILL_CHUNK=$ch timeout 120 python bench.py --strategy patch --steps 30 --warmup 3 --no-cpu-baseline --no-e2e --no-also 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']
    print('tile_bytes=$tb fill_chunk=$ch value=%.0f sweeps/s step=%.3f ms step_frac=%.3f kernels=%s' % (d['value'], d['ms_per_step'], r['step_frac'], {k: round(v,4) for k,v in r['kernels_ms_per_step'].items()}))
except Exception as e:
    print('tile_bytes=$tb fill_chunk=$ch FAILED', e)"
done
