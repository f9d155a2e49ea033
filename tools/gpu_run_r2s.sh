set -x
mkdir -p gpurun_out
which compute-sanitizer; compute-sanitizer --version | head -2
timeout 420 compute-sanitizer --tool memcheck --error-exitcode 3 --print-limit 20 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/memcheck_smoke_r2s.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/memcheck_smoke_r2s.log
tail -n 15 gpurun_out/memcheck_smoke_r2s.log
