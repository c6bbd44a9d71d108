# Round 2 (1 GPU): K3 DRAM traffic against the band size of the tile walk (QEMB_K3_GROUP) with the streaming stores,
# timing first (no profiler), then ncu dram bytes of one launch each.
#   gpurun --timeout 1500 -- 'bash scripts/gpu_calls/r2_call6.sh'
set -x
mkdir -p gpurun_out
for G in 8 12 16 24 32 48; do
  QEMB_K3_GROUP=$G timeout 120 python scripts/profile_k3.py 25425 2716 4 2>&1 | tail -n 1 | sed "s/^/group $G: /" | tee -a gpurun_out/r2c6_k3_group_timing.txt
done
for G in 12 16 24 32 48; do
  QEMB_K3_GROUP=$G timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct -k regex:gram_kernel --clock-control none -s 1 -c 1 --csv \
    --log-file gpurun_out/r2c6_k3_group_$G.csv python scripts/profile_k3.py 25425 2716 1 > /dev/null 2>&1; echo "ncu group $G exit $?"
  grep -E "dram__bytes|hit_rate|gpu__time" gpurun_out/r2c6_k3_group_$G.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' | sed "s/^/group $G: /" | tee -a gpurun_out/r2c6_k3_group_traffic.txt
done
# N = 1 end to end: fixed fractions of whole columns against the self-tuned balance (new AVX-512 tiled mirror)
timeout 600 python scripts/e2e_timeline.py --modes triangle@0:sharded,triangle@0.05:sharded,triangle@0.1:sharded,triangle@0.16:sharded,triangle@0.25:sharded,auto:sharded,full:sharded --reps 2 --out gpurun_out/r2c6_e2e_hybrid.jsonl > gpurun_out/r2c6_e2e_hybrid.log 2>&1; echo "hybrid exit $?"
python - <<'PY'
import json
for l in open('gpurun_out/r2c6_e2e_hybrid.jsonl'):
    d = json.loads(l); r = d['ranks'][0]
    print(d['output'], 'e2e', round(d['e2e_s_max_over_ranks'], 3), 'build', round(r['mean']['build_s'], 3), 'd2h_gb', round(r['mean']['d2h_gb'], 1), 'autotune', r.get('autotune'))
PY
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r2c6_pytest.log 2>&1; echo "pytest exit $?"; tail -n 5 gpurun_out/r2c6_pytest.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2c6_bench.json 2> gpurun_out/r2c6_bench.err; echo "bench exit $?"; tail -c 3000 gpurun_out/r2c6_bench.json; tail -n 3 gpurun_out/r2c6_bench.err
