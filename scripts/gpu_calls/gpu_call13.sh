set -x
mkdir -p gpurun_out
nproc; free -g | head -2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 4 --steps 3 --warmup 3 > gpurun_out/bench_r1l_n4.json 2> gpurun_out/bench_r1l_n4.err; echo "bench4 exit $?"
cat gpurun_out/bench_r1l_n4.json; tail -n 5 gpurun_out/bench_r1l_n4.err
