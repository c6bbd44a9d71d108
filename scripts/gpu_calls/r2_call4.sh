# Round 2 (2 GPUs): the multi-rank code paths under NCCL before the 8-GPU call -- the two 2-GPU tests, bench at N = 2
# (distributed staging, full-square output, e2e_consumer, sharded extra), e2e timeline at N = 2.
#   gpurun --gpus 2 --timeout 1500 -- 'bash scripts/gpu_calls/r2_call4.sh'
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests -m gpu -q -x  > gpurun_out/r2c4_pytest.log 2>&1; echo "pytest exit $?"; tail -n 8 gpurun_out/r2c4_pytest.log
timeout 900 $TR --nproc-per-node 2 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2c4_bench_n2.json 2> gpurun_out/r2c4_bench_n2.err; echo "bench n2 exit $?"; tail -c 4500 gpurun_out/r2c4_bench_n2.json; tail -n 5 gpurun_out/r2c4_bench_n2.err
timeout 600 $TR --nproc-per-node 2 --master-port 29512 scripts/e2e_timeline.py --modes triangle:replicated,triangle:sharded,full:sharded --reps 2 --out gpurun_out/e2e_timeline_r2.jsonl > gpurun_out/r2c4_timeline_n2.log 2>&1; echo "timeline n2 exit $?"; grep '^{' gpurun_out/r2c4_timeline_n2.log | cut -c1-900
