set -x
mkdir -p gpurun_out
echo "== 12 warps x 8 rows" > gpurun_out/jk_stream9.log
timeout 120 python scripts/profile_jk.py stream >> gpurun_out/jk_stream9.log 2>&1
echo "== 8 warps x 16 rows" >> gpurun_out/jk_stream9.log
QEMB_B200_LIB=$PWD/build_alt/libqemb_8x16.so timeout 120 python scripts/profile_jk.py stream >> gpurun_out/jk_stream9.log 2>&1
cat gpurun_out/jk_stream9.log
timeout 600 python -m pytest tests -m gpu -q -x -k "jk" > gpurun_out/pytest_gpu23.log 2>&1; echo "pytest exit $?"; tail -n 3 gpurun_out/pytest_gpu23.log
QEMB_B200_LIB=$PWD/build_alt/libqemb_8x16.so timeout 600 python -m pytest tests -m gpu -q -x -k "jk" > gpurun_out/pytest_gpu23b.log 2>&1; echo "pytest exit $?"; tail -n 3 gpurun_out/pytest_gpu23b.log
QEMB_B200_LIB=$PWD/build_alt/libqemb_8x16.so timeout 600 ncu --set full --clock-control none --import-source on -k regex:jk_stream -c 1 -f -o gpurun_out/jk_stream_r1w python scripts/profile_jk.py stream > gpurun_out/jk_ncu12.log 2>&1; echo "ncu exit $?"
