# First GPU call of round 2 (prepared at the end of round 1, when no GPU minutes were left): confirms on a B200
# what was changed after the last round-1 measurement and only checked through the restatement on the host --
# the rounded base-256 slicer (5 .. 8 slices), the NaN rule and the accumulating form of the int8-slice K3 -- then re-times the three settings.
#   gpurun --timeout 1500 -- 'bash scripts/gpu_calls/r2_call1.sh'
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/r2c1_pytest.log 2>&1; echo "pytest exit $?"; tail -n 6 gpurun_out/r2c1_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2c1_smoke.log 2>&1; echo "smoke exit $?"
for s in 0 5 6 7 8; do
  timeout 300 python bench.py --steps 3 --warmup 3 --k3-int8 $s > gpurun_out/r2c1_bench_int8x$s.json 2> gpurun_out/r2c1_bench_int8x$s.err; echo "bench $s exit $?"
  tail -c 600 gpurun_out/r2c1_bench_int8x$s.json
done
# launch list of the 7-slice run (kernel shares), after the plain runs above exited
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2c1_launches_int8x7.csv \
  python bench.py --steps 1 --warmup 3 --k3-int8 7 > gpurun_out/r2c1_ncu.log 2>&1; echo "ncu exit $?"
