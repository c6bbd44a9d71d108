set -x
mkdir -p gpurun_out
timeout 120 python scripts/profile_jk.py stream > gpurun_out/jk_stream12.log 2>&1; cat gpurun_out/jk_stream12.log
timeout 600 python -m pytest tests -m gpu -q -x -k "jk" > gpurun_out/pytest_gpu26.log 2>&1; echo "pytest exit $?"; tail -n 3 gpurun_out/pytest_gpu26.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jk_stream -c 1 -f -o gpurun_out/jk_stream_r1z python scripts/profile_jk.py stream > gpurun_out/jk_ncu15.log 2>&1; echo "ncu exit $?"
