set -x
mkdir -p gpurun_out
nvidia-smi -L; nproc
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
python bench.py > gpurun_out/bench_r1d.json 2> gpurun_out/bench_r1d.err; echo "bench exit $?"
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_ref_r1d.json 2> gpurun_out/bench_ref_r1d.err; echo "ref exit $?"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_s2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_r1d.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python scripts/profile_k3.py > gpurun_out/k3_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gram -s 1 -c 1 -o gpurun_out/k3_full_r1d python scripts/profile_k3.py > gpurun_out/k3_ncu.log 2>&1
python scripts/perf_probe.py c32 1 1 > gpurun_out/probe_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gather_gemm -s 2 -c 2 -o gpurun_out/gather_full_r1d python scripts/perf_probe.py c32 1 1 > gpurun_out/gather_ncu.log 2>&1
ls -la gpurun_out
tail -3 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/probe_plain.log
