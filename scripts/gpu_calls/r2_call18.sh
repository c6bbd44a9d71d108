# Round 2 (1 GPU), closing call: GPU suite with the caller-built-tensor test, K3 DRAM-traffic experiments (alternating k
# direction = QEMB_K3_HINT bit 2, L2 persisting set-aside = QEMB_L2_PERSIST_MB), short bench, --set full captures of K3
# and the gather GEMM for the round-2 record.
set -x
mkdir -p gpurun_out
timeout 420 python -m pytest tests -m gpu -q -x > gpurun_out/r2c18_pytest.log 2>&1; echo "pytest exit $?"; tail -n 4 gpurun_out/r2c18_pytest.log
QEMB_K3_HINT=5 timeout 150 python -m pytest tests -m gpu -q -x -k "steps_match or octane_full or medium_properties or accumulate_mode" > gpurun_out/r2c18_pytest_hint5.log 2>&1; echo "pytest hint5 exit $?"; tail -n 2 gpurun_out/r2c18_pytest_hint5.log
for cfg in "5 0" "1 40"; do set -- $cfg
  QEMB_K3_HINT=$1 QEMB_L2_PERSIST_MB=$2 timeout 60 python scripts/profile_k3.py 25425 2716 4 2>&1 | tail -n 2 | tr '\n' ' ' | sed "s/^/hint $1 persist $2: /" | tee -a gpurun_out/r2c18_k3_timing.txt; echo | tee -a gpurun_out/r2c18_k3_timing.txt
done
for cfg in "5 0" "1 40" "5 40"; do set -- $cfg
  QEMB_K3_HINT=$1 QEMB_L2_PERSIST_MB=$2 timeout 90 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct -k regex:gram_kernel --clock-control none -s 1 -c 1 --csv \
    --log-file gpurun_out/r2c18_k3_$1_$2.csv python scripts/profile_k3.py 25425 2716 1 > /dev/null 2>&1; echo "ncu hint $1 persist $2 exit $?"
  grep -E "dram__bytes|hit_rate|gpu__time" gpurun_out/r2c18_k3_$1_$2.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' | sed "s/^/hint $1 persist $2: /" | tee -a gpurun_out/r2c18_k3_traffic.txt
done
timeout 150 ncu --set full --clock-control none --import-source on -k regex:gram_kernel -s 1 -c 1 -f -o gpurun_out/k3_full_r2 python scripts/profile_k3.py 25425 2716 1 > gpurun_out/r2c18_k3_full.log 2>&1; echo "ncu full k3 exit $?"
timeout 150 ncu --set full --clock-control none --import-source on -k regex:gather_gemm -s 2 -c 2 -f -o gpurun_out/gather_full_r2 python scripts/perf_probe.py c32 1 1 > gpurun_out/r2c18_gather_full.log 2>&1; echo "ncu full gather exit $?"
timeout 240 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-file > gpurun_out/r2c18_bench.json 2> gpurun_out/r2c18_bench.err; echo "bench exit $?"; tail -c 300 gpurun_out/r2c18_bench.json
ls -la gpurun_out | tail -n 20
