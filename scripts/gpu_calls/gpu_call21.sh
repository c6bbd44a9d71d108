set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu21.log 2>&1; echo "pytest exit $?"; tail -n 6 gpurun_out/pytest_gpu21.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke21.log 2>&1; echo "smoke exit $?"; tail -n 3 gpurun_out/smoke21.log
timeout 200 python scripts/jk_sweep.py > gpurun_out/jk_sweep.log 2>&1; echo "jk_sweep exit $?"; cat gpurun_out/jk_sweep.log
timeout 900 python bench.py > gpurun_out/bench_r1p.json 2> gpurun_out/bench_r1p.err; echo "bench exit $?"; cut -c1-600 gpurun_out/bench_r1p.json
