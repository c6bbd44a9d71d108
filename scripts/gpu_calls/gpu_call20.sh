set -x
mkdir -p gpurun_out
timeout 120 python scripts/profile_jk.py both > gpurun_out/jk_stream7.log 2>&1; echo "profile exit $?"; tail -10 gpurun_out/jk_stream7.log
timeout 600 python -m pytest tests -m gpu -q -x -k "jk" > gpurun_out/pytest_gpu20.log 2>&1; echo "pytest exit $?"; tail -n 5 gpurun_out/pytest_gpu20.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jk_stream -c 1 -f -o gpurun_out/jk_stream_r1u python scripts/profile_jk.py stream > gpurun_out/jk_ncu10.log 2>&1; echo "ncu exit $?"
