# Round 2 (1 GPU), last minutes: the autotune tie-break (trials within 5 % -> rate model) on whatever box is drawn: its GPU
# test and a short bench.
set -x
mkdir -p gpurun_out
timeout 70 python -m pytest tests/test_gpu_seam.py -m gpu -q -x -k autotuned > gpurun_out/r2c19_pytest.log 2>&1; echo "pytest exit $?"; tail -n 2 gpurun_out/r2c19_pytest.log
timeout 100 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-file > gpurun_out/r2c19_bench.json 2> gpurun_out/r2c19_bench.err; echo "bench exit $?"
python -c "
import json; d=json.loads(open('gpurun_out/r2c19_bench.json').read().strip().splitlines()[-1]); e=d['e2e']
print('value',d['value'],'e2e',e['value'],'consumer',d['e2e_consumer']['value']); print(e['autotune_per_rank'], e['timeline_max_over_ranks'])"
