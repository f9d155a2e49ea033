#!/bin/bash
# Sweep the pipeline stage size (sweeps per sub-batch; 64 = no pipelining). bash tools/sub_batch_sweep.sh 2 4 8 ...
for sb in "$@"; do
  MV3D_SUB_BATCH=$sb python bench.py --steps 30 --warmup 3 --no-cpu-baseline --no-e2e --no-also 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('sub_batch=$sb value=%.0f sweeps/s step=%.3f ms step_frac=%.3f tile_frac=%.3f kernels=%s' % (d['value'], d['ms_per_step'], r['step_frac'], r['frac'], {k: round(v,4) for k,v in r['kernels_ms_per_step'].items()}))"
done
