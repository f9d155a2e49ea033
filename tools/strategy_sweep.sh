#!/bin/bash
# The reduction strategies side by side on bench.py's workload (config 2). Run on the GPU box: bash tools/strategy_sweep.sh tile atomic sort patch
for st in "$@"; do
  timeout 120 python bench.py --strategy $st --steps 30 --warmup 3 --no-cpu-baseline --no-e2e --no-also 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']
    print('strategy=$st value=%.0f sweeps/s step=%.3f ms step_frac=%.3f kernels=%s' % (d['value'], d['ms_per_step'], r['step_frac'], {k: round(v,4) for k,v in r['kernels_ms_per_step'].items()}))
except Exception as e:
    print('strategy=$st FAILED', e)"
done
