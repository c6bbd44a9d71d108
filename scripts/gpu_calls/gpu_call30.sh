set -x
mkdir -p gpurun_out
timeout 180 python scripts/profile_jk.py both > gpurun_out/jk_stream13.log 2>&1; cat gpurun_out/jk_stream13.log
timeout 600 python -m pytest tests -m gpu -q -x -k "jk" > gpurun_out/pytest_gpu30.log 2>&1; echo "pytest exit $?"; tail -n 5 gpurun_out/pytest_gpu30.log
QEMB_JK_KERNEL=stream12 timeout 600 ncu --set full --clock-control none --import-source on -k regex:jk_stream12 -c 1 -f -o gpurun_out/jk_stream12_r1 python scripts/profile_jk.py stream12 > gpurun_out/jk_ncu16.log 2>&1; echo "ncu exit $?"
