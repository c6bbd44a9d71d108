set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q -x -k "k3_int8" > gpurun_out/pytest_gpu40.log 2>&1; echo "pytest exit $?"; tail -n 6 gpurun_out/pytest_gpu40.log
