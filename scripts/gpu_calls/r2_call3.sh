# Round 2, call 2 (1 GPU): measured K3 DRAM traffic over the bench workload's own launches, launch list of one bench
# run, compute-sanitizer (memcheck / racecheck / synccheck) over every kernel on a golden case.
#   gpurun --timeout 2400 -- 'bash scripts/gpu_calls/r2_call3.sh'
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2c3_pytest.log 2>&1; echo "pytest exit $?"; tail -n 14 gpurun_out/r2c3_pytest.log
timeout 300 python scripts/k3_traffic_target.py > gpurun_out/r2c3_traffic_plain.log 2>&1; echo "target exit $?"
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:gram_kernel --clock-control none -c 40 --csv \
  --log-file gpurun_out/r2c3_k3_traffic.csv python scripts/k3_traffic_target.py > gpurun_out/r2c3_traffic_ncu.log 2>&1; echo "ncu traffic exit $?"
python scripts/k3_traffic_summarise.py gpurun_out/r2c3_k3_traffic.csv gpurun_out/k3_traffic_r2.json
timeout 300 python scripts/sanitize_target.py > gpurun_out/r2c3_sanitize_plain.log 2>&1; echo "sanitize target exit $?"; tail -n 2 gpurun_out/r2c3_sanitize_plain.log
for tool in memcheck synccheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 30 python scripts/sanitize_target.py > gpurun_out/r2c3_sanitizer_$tool.log 2>&1; echo "$tool exit $?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Error|done" gpurun_out/r2c3_sanitizer_$tool.log | head -12
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/r2c3_launches.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-file --e2e-steps 1 > gpurun_out/r2c3_ncu_bench.log 2>&1; echo "ncu launches exit $?"
timeout 300 python scripts/ref_cuda_bar.py cuda > gpurun_out/r2c3_refbar_shipped.log 2>&1; echo "ref bar shipped exit $?"; tail -n 3 gpurun_out/r2c3_refbar_shipped.log
timeout 300 python scripts/ref_cuda_bar.py cuda_betafix > gpurun_out/r2c3_refbar_betafix.log 2>&1; echo "ref bar betafix exit $?"; tail -n 3 gpurun_out/r2c3_refbar_betafix.log
timeout 600 python scripts/sweep.py --low --batch 32 --out gpurun_out/r2c3_sweep_low.json > gpurun_out/r2c3_sweep_low.log 2>&1; echo "sweep low exit $?"; tail -n 8 gpurun_out/r2c3_sweep_low.log | cut -c1-700
