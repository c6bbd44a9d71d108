set -x
mkdir -p gpurun_out
nvidia-smi -L; nproc; free -g | head -2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu11.log 2>&1; echo "pytest exit $?"
tail -n 6 gpurun_out/pytest_gpu11.log
timeout 900 $TR --master-port 29703 scripts/sharded_bench.py --case h2o64 --out gpurun_out/sharded_r1j.json > gpurun_out/sharded_h2o64.log 2>&1; echo "h2o64 exit $?"
tail -n 4 gpurun_out/sharded_h2o64.log
timeout 600 $TR --master-port 29704 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_r1j_n2.json 2> gpurun_out/bench_r1j_n2.err; echo "bench2 exit $?"
cat gpurun_out/bench_r1j_n2.json
