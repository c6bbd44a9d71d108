# Round 2 (8 GPUs): where the end-to-end time goes at N = 8 / 4 (old and new staging / output modes), host DRAM
# bandwidth at several thread counts, bench at N = 8 with the sharded extra.
#   gpurun --gpus 8 --timeout 1200 -- 'bash scripts/gpu_calls/r2_call5.sh'
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nproc; free -g | head -2; nvidia-smi topo -m | head -12
python scripts/host_bw_probe.py 4 8 16 32 > gpurun_out/host_bw_r2_8gpu_box.json 2>&1; cat gpurun_out/host_bw_r2_8gpu_box.json
timeout 420 $TR --nproc-per-node 8 --master-port 29521 scripts/e2e_timeline.py --modes triangle:replicated,triangle:sharded,full:sharded --reps 2 --out gpurun_out/e2e_timeline_r2.jsonl > gpurun_out/r2c5_timeline_n8.log 2>&1; echo "timeline n8 exit $?"; grep '^{' gpurun_out/r2c5_timeline_n8.log | cut -c1-900
timeout 300 $TR --nproc-per-node 4 --master-port 29522 scripts/e2e_timeline.py --modes triangle:sharded,full:sharded --reps 2 --out gpurun_out/e2e_timeline_r2.jsonl > gpurun_out/r2c5_timeline_n4.log 2>&1; echo "timeline n4 exit $?"; grep '^{' gpurun_out/r2c5_timeline_n4.log | cut -c1-900
timeout 600 $TR --nproc-per-node 8 --master-port 29523 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2c5_bench_n8.json 2> gpurun_out/r2c5_bench_n8.err; echo "bench n8 exit $?"; tail -c 4500 gpurun_out/r2c5_bench_n8.json; tail -n 5 gpurun_out/r2c5_bench_n8.err
