# Round 2 (1 GPU): K3 L2 eviction hints on the TMA operand loads x band size (timing, then ncu DRAM bytes), full GPU
# suite, bench N = 1.
#   gpurun --timeout 1800 -- 'bash scripts/gpu_calls/r2_call8.sh'
set -x
mkdir -p gpurun_out
for cfg in "0 12" "1 12" "2 12" "1 16" "2 16" "1 24" "2 24" "2 32"; do set -- $cfg
  QEMB_K3_HINT=$1 QEMB_K3_GROUP=$2 timeout 120 python scripts/profile_k3.py 25425 2716 4 2>&1 | tail -n 1 | sed "s/^/hint $1 group $2: /" | tee -a gpurun_out/r2c8_k3_hint_timing.txt
done
for cfg in "1 12" "2 12" "1 16" "2 16" "1 24" "2 24" "2 32"; do set -- $cfg
  QEMB_K3_HINT=$1 QEMB_K3_GROUP=$2 timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct -k regex:gram_kernel --clock-control none -s 1 -c 1 --csv \
    --log-file gpurun_out/r2c8_k3_hint_$1_$2.csv python scripts/profile_k3.py 25425 2716 1 > /dev/null 2>&1; echo "ncu hint $1 group $2 exit $?"
  grep -E "dram__bytes|hit_rate|gpu__time" gpurun_out/r2c8_k3_hint_$1_$2.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' | sed "s/^/hint $1 group $2: /" | tee -a gpurun_out/r2c8_k3_hint_traffic.txt
done
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r2c8_pytest.log 2>&1; echo "pytest exit $?"; tail -n 5 gpurun_out/r2c8_pytest.log
