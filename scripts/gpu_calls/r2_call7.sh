# Round 2 (8 GPUs), final delivery check: self-tuned PCIe / host-mirror split with and without the cross-rank fragment
# queue against full squares, at N = 8 and N = 4; then the bench line at N = 8.
#   gpurun --gpus 8 --timeout 1200 -- 'bash scripts/gpu_calls/r2_call7.sh'
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 420 $TR --nproc-per-node 8 --master-port 29531 scripts/e2e_timeline.py --modes auto:sharded:dynamic,auto:sharded,full:sharded:dynamic --reps 3 --out gpurun_out/e2e_timeline_r2_final.jsonl > gpurun_out/r2c7_timeline_n8.log 2>&1; echo "timeline n8 exit $?"; grep '^{' gpurun_out/r2c7_timeline_n8.log | cut -c1-700; tail -n 3 gpurun_out/r2c7_timeline_n8.log | cut -c1-300
timeout 300 $TR --nproc-per-node 4 --master-port 29532 scripts/e2e_timeline.py --modes auto:sharded:dynamic,full:sharded:dynamic --reps 2 --out gpurun_out/e2e_timeline_r2_final.jsonl > gpurun_out/r2c7_timeline_n4.log 2>&1; echo "timeline n4 exit $?"; grep '^{' gpurun_out/r2c7_timeline_n4.log | cut -c1-700
timeout 600 $TR --nproc-per-node 8 --master-port 29533 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2c7_bench_n8.json 2> gpurun_out/r2c7_bench_n8.err; echo "bench n8 exit $?"; tail -c 6000 gpurun_out/r2c7_bench_n8.json | cut -c1-6000; tail -n 5 gpurun_out/r2c7_bench_n8.err
python - <<'PY'
import json
for l in open('gpurun_out/e2e_timeline_r2_final.jsonl'):
    d = json.loads(l)
    print(d['n_gpus'], d['output'], d['upload'], 'dynamic' if d.get('dynamic_queue') else 'static', 'e2e', round(d['e2e_s_max_over_ranks'], 3))
    for r in d['ranks']:
        m = r['mean']
        print('   rank', r['rank'], 'frags', m.get('fragments_done'), 'build', round(m['build_s'], 3), 'd2h_gb', round(m['d2h_gb'], 1), 'tune', {k: round(v, 2) for k, v in (r.get('autotune') or {}).items()})
PY
