# Round 2 (2 GPUs): the final code at N = 2 the way the driver launches it (both arms), plus the two 2-GPU tests.
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 $TR --nproc-per-node 2 --master-port 29541 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2c12_bench_n2.json 2> gpurun_out/r2c12_bench_n2.err; echo "bench n2 exit $?"
python -c "
import json; d=json.loads(open('gpurun_out/r2c12_bench_n2.json').read().strip().splitlines()[-1]); e=d['e2e']
print('value',d['value'],'e2e',e['value'],'consumer',d['e2e_consumer']['value']); print(e['autotune_per_rank'], e['fragments_taken_per_rank_last_step'], e['timeline_max_over_ranks'], e['d2h_roofline_frac']); print(d['sharded']['parity_vs_oracle'], d['sharded']['oversize']['seconds_per_fragment'], d['sharded']['oversize']['gen_s'])"
tail -n 3 gpurun_out/r2c12_bench_n2.err
