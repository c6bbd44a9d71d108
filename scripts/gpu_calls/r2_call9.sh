# Round 2 (1 GPU), closing call: full GPU suite, smoke, K3 DRAM traffic over the workload's own launches with the final
# kernel, the bench line with driver-like flags.
#   gpurun --timeout 2400 -- 'bash scripts/gpu_calls/r2_call9.sh'
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/r2c9_pytest.log 2>&1; echo "pytest exit $?"; tail -n 12 gpurun_out/r2c9_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2c9_smoke.log 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/r2c9_smoke.log
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:gram_kernel --clock-control none -c 40 --csv \
  --log-file gpurun_out/r2c9_k3_traffic.csv python scripts/k3_traffic_target.py > gpurun_out/r2c9_traffic_ncu.log 2>&1; echo "ncu traffic exit $?"
python scripts/k3_traffic_summarise.py gpurun_out/r2c9_k3_traffic.csv gpurun_out/k3_traffic_r2.json
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2c9_bench.json 2> gpurun_out/r2c9_bench.err; echo "bench exit $?"; tail -c 2500 gpurun_out/r2c9_bench.json; tail -n 3 gpurun_out/r2c9_bench.err
