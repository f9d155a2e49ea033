#!/bin/bash
# Overlapped ATOMIC strategy: sweeps per fill chunk x in-flight bulk stores. bash tools/atomic_sweep.sh "4 8 16" "4 32"
for fc in $1; do for fp in ${2:-4}; do
  MV3D_ATOMIC_OVERLAP=1 MV3D_FILL_CHUNK=$fc MV3D_FILL_PENDING=$fp python bench.py --strategy atomic --steps 30 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('fill_chunk=$fc pending=$fp value=%.0f sweeps/s step=%.3f ms step_frac=%.3f kernels=%s' % (d['value'], d['ms_per_step'], r['step_frac'], {k: round(v,4) for k,v in r['kernels_ms_per_step'].items()}))"
done; done
