set -x
mkdir -p gpurun_out
timeout 300 python scripts/profile_jk.py both > gpurun_out/jk_stream1.log 2>&1; echo "profile exit $?"; cat gpurun_out/jk_stream1.log | tail -12
timeout 600 python -m pytest tests -m gpu -q -x -k "jk" > gpurun_out/pytest_gpu18.log 2>&1; echo "pytest exit $?"; tail -n 12 gpurun_out/pytest_gpu18.log
