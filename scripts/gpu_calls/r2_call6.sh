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
# N = 1 end to end: how much of the upper triangle should cross PCIe instead of being mirrored by the host threads
for F in 0 0.1 0.16 0.22 0.3; do
  QEMB_FULL_COLUMN_FRACTION=$F timeout 300 python scripts/e2e_timeline.py --modes triangle:sharded --reps 2 --out gpurun_out/r2c6_e2e_hybrid.jsonl 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('full_frac $F', 'e2e', round(d['e2e_s_max_over_ranks'],3), 'build', round(d['rank0_mean']['build_s'],3), 'd2h_gb', round(d['rank0_mean']['d2h_gb'],1))" | tee -a gpurun_out/r2c6_e2e_hybrid.txt
done
