set -x
mkdir -p gpurun_out
QEMB_B200_LIB=$PWD/build_alt/libqemb_nocompute.so timeout 120 python scripts/profile_jk.py stream > gpurun_out/jk_nocompute.log 2>&1; tail -3 gpurun_out/jk_nocompute.log
