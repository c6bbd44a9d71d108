set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "k3_int8 or golden or smoke" > gpurun_out/pytest_gpu37.log 2>&1; echo "pytest exit $?"; tail -n 12 gpurun_out/pytest_gpu37.log
timeout 600 python bench.py --k3-int8 7 --no-cpu-baseline > gpurun_out/bench_r1r_int8x7.json 2> gpurun_out/bench_r1r_int8x7.err; echo "bench exit $?"; python -c "
import json;d=json.loads(open('gpurun_out/bench_r1r_int8x7.json').read().strip().splitlines()[-1]);print(d['value'],d['e2e']['value'],d['per_kernel'],d['config']['k3'])"; tail -3 gpurun_out/bench_r1r_int8x7.err
