set -x
mkdir -p gpurun_out
for w in 5 6 7 8 16; do echo "== JS_WSM $w"; QEMB_B200_LIB=$PWD/build_alt/libqemb_wsm$w.so timeout 120 python scripts/profile_jk.py stream 2>&1 | tail -2; done > gpurun_out/jk_wsm.log 2>&1
cat gpurun_out/jk_wsm.log
