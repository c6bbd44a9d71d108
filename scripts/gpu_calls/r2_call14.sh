# Round 2 (1 GPU), last check of the final tree: full GPU suite, smoke, bench with default flags.
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2c14_pytest.log 2>&1; echo "pytest exit $?"; tail -n 4 gpurun_out/r2c14_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2c14_smoke.log 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/r2c14_smoke.log
timeout 600 python bench.py > gpurun_out/r2c14_bench.json 2> gpurun_out/r2c14_bench.err; echo "bench exit $?"; python -c "
import json; d=json.loads(open('gpurun_out/r2c14_bench.json').read().strip().splitlines()[-1]); e=d['e2e']
print('value',d['value'],'e2e',e['value'],'consumer',d['e2e_consumer']['value'],'frac',d['roofline']['frac'],'traffic',d['roofline']['traffic']); print(e['autotune_per_rank']); print(d['k3_int8_extra']); print(d['cpu_baseline']['parity_per_fragment'])"
