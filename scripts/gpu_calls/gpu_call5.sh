set -x
mkdir -p gpurun_out
timeout 600 python scripts/diag_parity.py octane > gpurun_out/diag_octane.log 2>&1; echo "diag octane exit $?"
timeout 600 python scripts/diag_parity.py c32 > gpurun_out/diag_c32.log 2>&1; echo "diag c32 exit $?"
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu5.log 2>&1; echo "pytest exit $?"
timeout 1500 python scripts/sweep.py --out gpurun_out/sweep_r1e.json > gpurun_out/sweep.log 2>&1; echo "sweep exit $?"
cat gpurun_out/diag_octane.log gpurun_out/diag_c32.log; tail -n 30 gpurun_out/pytest_gpu5.log; tail -n 5 gpurun_out/sweep.log
