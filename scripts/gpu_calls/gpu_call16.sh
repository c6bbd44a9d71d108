set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -k "hybrid or pipelined" > gpurun_out/pytest_gpu16.log 2>&1; echo "pytest exit $?"; tail -n 8 gpurun_out/pytest_gpu16.log
for f in 0.0 0.05 0.08 0.12 0.18; do timeout 300 python scripts/e2e_timeline.py 1 3 32 $f 2>&1 | grep "^rep [12]"; done > gpurun_out/e2e_hybrid.log
cat gpurun_out/e2e_hybrid.log
