# Round 2, call 1 (1 GPU): everything changed on the host side so far -- seam / staging / writer / consumer pipeline /
# baseline-parity tests, smoke, the rewritten bench (both arms).
#   gpurun --timeout 2400 -- 'bash scripts/gpu_calls/r2_call2.sh'
set -x
mkdir -p gpurun_out
nproc; free -g | head -2; df -h /tmp | tail -1
timeout 1500 python -m pytest tests -m gpu -q -x --durations=15 > gpurun_out/r2c2_pytest.log 2>&1; echo "pytest exit $?"; tail -n 30 gpurun_out/r2c2_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2c2_smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/r2c2_smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2c2_bench.json 2> gpurun_out/r2c2_bench.err; echo "bench exit $?"; tail -c 5000 gpurun_out/r2c2_bench.json; tail -n 5 gpurun_out/r2c2_bench.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2c2_ref.json 2> gpurun_out/r2c2_ref.err; echo "ref exit $?"; tail -c 3000 gpurun_out/r2c2_ref.json; tail -n 5 gpurun_out/r2c2_ref.err
