# Round 2 (1 GPU): N = 1 delivery against the copy-group size of the triangle transfer (does an earlier mirror find the
# data in the last-level cache?)
set -x
mkdir -p gpurun_out
timeout 600 python scripts/e2e_timeline.py --modes triangle@0@64:sharded,triangle@0@8:sharded,triangle@0@16:sharded,triangle@0@32:sharded,triangle@0@128:sharded,triangle@0@256:sharded --reps 2 --out gpurun_out/r2c15_groups.jsonl > gpurun_out/r2c15_groups.log 2>&1; echo "exit $?"
python - <<'PY'
import json
for l in open('gpurun_out/r2c15_groups.jsonl'):
    d = json.loads(l); m = d['ranks'][0]['mean']
    print(d['output'], 'e2e', round(d['e2e_s_max_over_ranks'], 3), 'build', round(m['build_s'], 3), 'wait', round(m['wait_s'], 3))
PY
