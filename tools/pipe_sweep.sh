#!/bin/bash
# Sweep the pipelined tile kernel (MV3D_TILE_PIPE=G,threads,chunk,pending [@tile_bytes]); prints sweeps/s and the tile
# kernel's time per launch for each setting. Run on the GPU box:  bash tools/pipe_sweep.sh 0 5,192,8640,0 4,224,8640,0@45000
for cfg in "$@"; do
  pipe=${cfg%@*}; tb=36864
  if [[ "$cfg" == *@* ]]; then tb=${cfg#*@}; fi
  MV3D_TILE_PIPE=$pipe MV3D_TILE_BYTES=$tb timeout 120 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --no-e2e --no-also 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']
    print('pipe=$pipe tile_bytes=$tb value=%.0f sweeps/s step=%.3f ms tile_ms=%.4f frac=%.3f kernels=%s' % (d['value'], d['ms_per_step'], r['kernels_ms_per_step'].get('tile_build_kernel',0), r['frac'], {k: round(v,4) for k,v in r['kernels_ms_per_step'].items()}))
except Exception as e:
    print('pipe=$pipe tile_bytes=$tb FAILED', e)"
done
