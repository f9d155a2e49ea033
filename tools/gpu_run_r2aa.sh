set -x
mkdir -p gpurun_out
: > gpurun_out/variants_r2aa.txt
VARIANT_ENV="tile run" timeout 900 bash tools/variant_bench.sh "" "-DMV3D_TILE_RUN=2" "-DMV3D_TILE_RUN=4" "-DMV3D_TILE_RUN=8" "-DMV3D_TILE_RUN=16" >> gpurun_out/variants_r2aa.txt 2>&1
cat gpurun_out/variants_r2aa.txt
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "batch or sweeps or knobs" 2>&1 | tail -n 2
python -c "from mv3d_b200 import _lib; _lib.build_library(force=True)"
