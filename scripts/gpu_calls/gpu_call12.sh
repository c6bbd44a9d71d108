set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29712 scripts/sharded_bench.py --case chain400 --fused --out gpurun_out/sharded_r1k.json > gpurun_out/fused_chain400.log 2>&1; echo "chain400 exit $?"; tail -n 5 gpurun_out/fused_chain400.log
