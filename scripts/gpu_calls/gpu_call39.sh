set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "k3_int8 or golden or octane or stress" > gpurun_out/pytest_gpu39.log 2>&1; echo "pytest exit $?"; tail -n 4 gpurun_out/pytest_gpu39.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --k3-int8 8 --no-cpu-baseline > gpurun_out/bench_r1s_int8x8.json 2> gpurun_out/bench_r1s_int8x8.err; echo "bench exit $?"; python -c "
import json;d=json.loads(open('gpurun_out/bench_r1s_int8x8.json').read().strip().splitlines()[-1]);print(d['value'],d['e2e']['value'],d['per_kernel']['K3'])"
