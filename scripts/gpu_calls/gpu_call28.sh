set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "jk or smoke or abi" > gpurun_out/pytest_gpu28.log 2>&1; echo "pytest exit $?"; tail -n 3 gpurun_out/pytest_gpu28.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 120 python scripts/profile_jk.py stream 2>&1 | tail -2
