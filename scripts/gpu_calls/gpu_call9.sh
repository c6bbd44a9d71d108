set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu9.log 2>&1; echo "pytest exit $?"
tail -n 12 gpurun_out/pytest_gpu9.log
timeout 300 python scripts/profile_jk.py > gpurun_out/jk_plain2.log 2>&1; cat gpurun_out/jk_plain2.log
timeout 900 python scripts/ref_cuda_bar.py > gpurun_out/ref_cuda_bar.log 2>&1; echo "bar exit $?"; tail -n 6 gpurun_out/ref_cuda_bar.log
ncu --set full --clock-control none --import-source on -k regex:jk_rows -s 1 -c 1 -o gpurun_out/jk_full_r1h python scripts/profile_jk.py > gpurun_out/jk_ncu2.log 2>&1
