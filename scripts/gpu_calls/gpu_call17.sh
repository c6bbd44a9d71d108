set -x
mkdir -p gpurun_out
nproc; free -g | head -2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/bench_r1n_n8.json 2> gpurun_out/bench_r1n_n8.err; echo "bench8 exit $?"
cat gpurun_out/bench_r1n_n8.json; tail -n 5 gpurun_out/bench_r1n_n8.err
