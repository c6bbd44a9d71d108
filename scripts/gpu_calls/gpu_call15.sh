set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -k "sharded or fused" > gpurun_out/pytest_gpu15.log 2>&1; echo "pytest exit $?"
tail -n 15 gpurun_out/pytest_gpu15.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29731 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_r1m_n2.json 2> gpurun_out/bench_r1m_n2.err; echo "bench2 exit $?"
cat gpurun_out/bench_r1m_n2.json | cut -c1-400; tail -n 3 gpurun_out/bench_r1m_n2.err
