set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "knobs" > gpurun_out/pytest_knobs_r2m.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_knobs_r2m.log
tail -n 3 gpurun_out/pytest_knobs_r2m.log
: > gpurun_out/variants_r2m.txt
for ov in 0 32 16 8; do
VARIANT_ENV="OVERLAP=$ov prefetch" MV3D_OVERLAP=$ov timeout 300 bash tools/variant_bench.sh "" >> gpurun_out/variants_r2m.txt 2>&1
done
VARIANT_ENV="OVERLAP=16 noprefetch" MV3D_OVERLAP=16 MV3D_OVERLAP_PREFETCH=0 timeout 300 bash tools/variant_bench.sh "" >> gpurun_out/variants_r2m.txt 2>&1
cat gpurun_out/variants_r2m.txt
