set -x
mkdir -p gpurun_out
nvidia-smi -L; nproc; free -g | head -2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest exit $?"
timeout 300 $TR --master-port 29701 scripts/sharded_bench.py --case tiny --out gpurun_out/sharded_r1e.json > gpurun_out/sharded_tiny.log 2>&1; echo "tiny exit $?"
timeout 900 $TR --master-port 29702 scripts/sharded_bench.py --case chain400 --out gpurun_out/sharded_r1e.json > gpurun_out/sharded_chain400.log 2>&1; echo "chain400 exit $?"
timeout 900 $TR --master-port 29703 scripts/sharded_bench.py --case h2o64 --out gpurun_out/sharded_r1e.json > gpurun_out/sharded_h2o64.log 2>&1; echo "h2o64 exit $?"
timeout 600 $TR --master-port 29704 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_r1e_n2.json 2> gpurun_out/bench_r1e_n2.err; echo "bench2 exit $?"
tail -n 3 gpurun_out/pytest_gpu2.log gpurun_out/sharded_*.log
cat gpurun_out/sharded_r1e.json gpurun_out/bench_r1e_n2.json
