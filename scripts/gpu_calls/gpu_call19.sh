set -x
mkdir -p gpurun_out
./scripts/micro/dfma_rate > gpurun_out/dfma_rate.txt 2>&1; cat gpurun_out/dfma_rate.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jk_stream -c 1 -o gpurun_out/jk_stream_r1o python scripts/profile_jk.py stream > gpurun_out/jk_ncu4.log 2>&1; echo "ncu exit $?"; tail -3 gpurun_out/jk_ncu4.log
