set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "knobs" > gpurun_out/pytest_knobs_r2l.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_knobs_r2l.log
tail -n 5 gpurun_out/pytest_knobs_r2l.log
: > gpurun_out/variants_r2l.txt
for ov in 0 32 16 8 4; do
VARIANT_ENV="OVERLAP=$ov" MV3D_OVERLAP=$ov timeout 300 bash tools/variant_bench.sh "" >> gpurun_out/variants_r2l.txt 2>&1
done
cat gpurun_out/variants_r2l.txt
