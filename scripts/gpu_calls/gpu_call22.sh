set -x
mkdir -p gpurun_out
timeout 120 python scripts/profile_jk.py stream > gpurun_out/jk_stream8.log 2>&1; echo "profile exit $?"; tail -4 gpurun_out/jk_stream8.log
timeout 600 python -m pytest tests -m gpu -q -x -k "jk" > gpurun_out/pytest_gpu22.log 2>&1; echo "pytest exit $?"; tail -n 5 gpurun_out/pytest_gpu22.log
timeout 200 python scripts/jk_sweep.py --out gpurun_out/jk_sweep2.json > gpurun_out/jk_sweep2.log 2>&1; cut -c1-120 gpurun_out/jk_sweep2.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jk_stream -c 1 -f -o gpurun_out/jk_stream_r1v python scripts/profile_jk.py stream > gpurun_out/jk_ncu11.log 2>&1; echo "ncu exit $?"
