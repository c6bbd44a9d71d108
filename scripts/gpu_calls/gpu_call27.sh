set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_final_n2.log 2>&1; echo "pytest exit $?"; tail -n 5 gpurun_out/pytest_gpu_final_n2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29751 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_r1q_n2.json 2> gpurun_out/bench_r1q_n2.err; echo "bench2 exit $?"; python -c "
import json;d=json.loads(open('gpurun_out/bench_r1q_n2.json').read().strip().splitlines()[-1]);print(d['value'],d['e2e']['value'],d['roofline']['frac'],d['consumer'])"
