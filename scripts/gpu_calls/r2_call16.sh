# Round 2 (1 GPU): after raising the default copy-group size: output-path tests and a short bench.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "hybrid or pinned or autotuned or lanes or seam or stress or octane" > gpurun_out/r2c16_pytest.log 2>&1; echo "pytest exit $?"; tail -n 3 gpurun_out/r2c16_pytest.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-file > gpurun_out/r2c16_bench.json 2> gpurun_out/r2c16_bench.err; echo "bench exit $?"; python -c "
import json; d=json.loads(open('gpurun_out/r2c16_bench.json').read().strip().splitlines()[-1]); e=d['e2e']
print('value',d['value'],'e2e',e['value'],'consumer',d['e2e_consumer']['value']); print(e['autotune_per_rank'], e['timeline_max_over_ranks'])"
