#!/bin/bash
# Sweep the PATCH strategy's knobs. bash tools/patch_sweep.sh "<fill chunks>" "<tile bytes>" "<pending>" "<fill ctas>"
for fc in $1; do for tb in $2; do for fp in ${3:-4}; do for fa in ${4:-2}; do
  MV3D_FILL_CHUNK=$fc MV3D_PATCH_TILE_BYTES=$tb MV3D_FILL_PENDING=$fp MV3D_FILL_CTAS=$fa python bench.py --strategy patch --steps 30 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('fill_chunk=$fc tile_bytes=$tb pending=$fp ctas=$fa value=%.0f sweeps/s step=%.3f ms step_frac=%.3f kernels=%s' % (d['value'], d['ms_per_step'], r['step_frac'], {k: round(v,4) for k,v in r['kernels_ms_per_step'].items()}))"
done; done; done; done
