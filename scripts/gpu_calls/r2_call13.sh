# Round 2 (2 GPUs): the empirical autotune (trial deliveries) at N = 2 and, on the same box, at N = 1; autotune test.
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests -m gpu -q -x -k "autotuned or hybrid or pinned_ring or lanes" > gpurun_out/r2c13_pytest.log 2>&1; echo "pytest exit $?"; tail -n 3 gpurun_out/r2c13_pytest.log
timeout 600 $TR --nproc-per-node 2 --master-port 29551 bench.py --gpus 2 --steps 5 --warmup 3 --no-sharded-extra > gpurun_out/r2c13_bench_n2.json 2> gpurun_out/r2c13_bench_n2.err; echo "bench n2 exit $?"
python -c "
import json; d=json.loads(open('gpurun_out/r2c13_bench_n2.json').read().strip().splitlines()[-1]); e=d['e2e']
print('N=2 value',d['value'],'e2e',e['value'],'consumer',d['e2e_consumer']['value']); print(e['autotune_per_rank'], e['timeline_max_over_ranks'])"
tail -n 3 gpurun_out/r2c13_bench_n2.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-file > gpurun_out/r2c13_bench_n1.json 2> gpurun_out/r2c13_bench_n1.err; echo "bench n1 exit $?"
python -c "
import json; d=json.loads(open('gpurun_out/r2c13_bench_n1.json').read().strip().splitlines()[-1]); e=d['e2e']
print('N=1 value',d['value'],'e2e',e['value'],'consumer',d['e2e_consumer']['value']); print(e['autotune_per_rank'], e['timeline_max_over_ranks'])"
tail -n 3 gpurun_out/r2c13_bench_n1.err
