#!/bin/bash
# Sweep the tile size (bytes of output per shared-memory tile) of the tile strategy; prints sweeps/s and the
# tile kernel's time per launch for each size. Run on the GPU box:  bash tools/tile_sweep.sh 12288 16384 ...
for tb in "$@"; do
  MV3D_TILE_BYTES=$tb python bench.py --steps 30 --warmup 3 --no-cpu-baseline --no-e2e --no-also 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('tile_bytes=$tb value=%.0f sweeps/s step=%.3f ms tile_ms=%.4f frac=%.3f kernels=%s' % (d['value'], d['ms_per_step'], r['kernels_ms_per_step'].get('tile_build_kernel',0), r['frac'], {k: round(v,4) for k,v in r['kernels_ms_per_step'].items()}))"
done
