set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "jk or veff" > gpurun_out/pytest_gpu10.log 2>&1; echo "pytest exit $?"
tail -n 15 gpurun_out/pytest_gpu10.log
timeout 300 python scripts/profile_jk.py > gpurun_out/jk_plain3.log 2>&1; cat gpurun_out/jk_plain3.log
ncu --set full --clock-control none --import-source on -k regex:jk_rows -s 1 -c 1 -o gpurun_out/jk_full_r1i python scripts/profile_jk.py > gpurun_out/jk_ncu3.log 2>&1
