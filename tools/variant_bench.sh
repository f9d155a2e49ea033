#!/bin/bash
# Rebuild the library with extra nvcc flags and time config 2 for each variant (run on the GPU box):
#   bash tools/variant_bench.sh "" "-DMV3D_CLASSIFY_MIN_BLOCKS=6" "-DMV3D_POINT_ITEMS=4"
# Environment knobs (MV3D_TILE_BYTES=...) are inherited. The last variant stays built.
for extra in "$@"; do
  MV3D_NVCC_EXTRA="$extra" python -c "from mv3d_b200 import _lib; _lib.build_library(force=True)" || { echo "variant '$extra' BUILD FAILED"; continue; }
  timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --no-e2e --no-also 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']
    print('variant=[$extra] env=[${VARIANT_ENV}] value=%.0f sweeps/s step=%.4f ms step_frac=%.3f frac=%.3f kernels=%s' % (d['value'], d['ms_per_step'], r['step_frac'], r['frac'], {k: round(v,4) for k,v in r['kernels_ms_per_step'].items()}))
except Exception as e:
    print('variant=[$extra] FAILED', e)"
done
