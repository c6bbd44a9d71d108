set -x
mkdir -p gpurun_out
g++ -O3 -pthread -Iquemb_b200/csrc scripts/host_bw.cpp quemb_b200/csrc/host_pool.cpp -o /tmp/hbw && /tmp/hbw > gpurun_out/host_bw.log 2>&1
g++ -O3 -pthread -DRBV=32 -DCBV=128 -DQEMB_MIRROR_RB=32 -DQEMB_MIRROR_CB=128 -Iquemb_b200/csrc scripts/host_bw.cpp quemb_b200/csrc/host_pool.cpp -o /tmp/hbw2 && /tmp/hbw2 | tail -2 >> gpurun_out/host_bw.log 2>&1
lscpu | head -25 > gpurun_out/lscpu.log; numactl -H >> gpurun_out/lscpu.log 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
python scripts/e2e_timeline.py 1 3 > gpurun_out/tl_tri3.log 2>&1
python scripts/e2e_timeline.py 0 3 > gpurun_out/tl_full3.log 2>&1
python scripts/e2e_timeline.py 0 2 > gpurun_out/tl_full2.log 2>&1
tail -5 gpurun_out/tl_*.log gpurun_out/host_bw.log
