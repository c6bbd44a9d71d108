set -x
mkdir -p gpurun_out
timeout 1200 python scripts/race_check.py 25 > gpurun_out/race_check.log 2>&1; echo "race exit $?"
cat gpurun_out/race_check.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu7.log 2>&1; echo "pytest exit $?"
tail -n 5 gpurun_out/pytest_gpu7.log
python bench.py > gpurun_out/bench_r1f.json 2> gpurun_out/bench_r1f.err; echo "bench exit $?"
cat gpurun_out/bench_r1f.json
