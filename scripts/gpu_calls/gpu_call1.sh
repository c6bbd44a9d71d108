set -x
nvidia-smi -L
nproc; free -g | head -2
nvcc -O2 -o /tmp/pcie_bw scripts/pcie_bw.cu && /tmp/pcie_bw > gpurun_out/pcie.log 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
python bench.py > gpurun_out/bench_r1b.json 2> gpurun_out/bench_r1b.err; echo "bench exit $?"
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_ref_r1b.json 2> gpurun_out/bench_ref_r1b.err; echo "ref exit $?"
python scripts/profile_k3.py > gpurun_out/k3_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gram_kernel -s 1 -c 2 -o gpurun_out/k3_full python scripts/profile_k3.py > gpurun_out/k3_ncu.log 2>&1
python scripts/perf_probe.py c32 1 1 > gpurun_out/probe_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gather_gemm -s 2 -c 2 -o gpurun_out/gather_full python scripts/perf_probe.py c32 1 1 > gpurun_out/gather_ncu.log 2>&1
ls -la gpurun_out
