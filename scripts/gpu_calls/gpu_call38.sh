set -x
mkdir -p gpurun_out
timeout 600 python bench.py --k3-int8 7 > gpurun_out/bench_r1s_int8x7.json 2> gpurun_out/bench_r1s_int8x7.err; echo "bench exit $?"; python -c "
import json;d=json.loads(open('gpurun_out/bench_r1s_int8x7.json').read().strip().splitlines()[-1]);print(d['value'],d['e2e']['value'],d['cpu_baseline'])"
