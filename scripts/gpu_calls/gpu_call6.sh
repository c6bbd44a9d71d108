set -x
mkdir -p gpurun_out
timeout 900 python scripts/diag2.py octane > gpurun_out/diag2_octane.log 2>&1; echo "exit $?"
timeout 900 python scripts/diag2.py c32 > gpurun_out/diag2_c32.log 2>&1; echo "exit $?"
cat gpurun_out/diag2_octane.log gpurun_out/diag2_c32.log
