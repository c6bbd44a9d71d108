set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu8.log 2>&1; echo "pytest exit $?"
tail -n 5 gpurun_out/pytest_gpu8.log
timeout 1500 python scripts/sweep.py --out gpurun_out/sweep_r1g.json > gpurun_out/sweep_g.log 2>&1; echo "sweep exit $?"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_s2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_r1g.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python scripts/profile_k3.py > gpurun_out/k3_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gram_kernel -s 1 -c 1 -o gpurun_out/k3_full_r1g python scripts/profile_k3.py > gpurun_out/k3_ncu.log 2>&1
python scripts/perf_probe.py c32 1 1 > gpurun_out/probe_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gather_gemm -s 2 -c 2 -o gpurun_out/gather_full_r1g python scripts/perf_probe.py c32 1 1 > gpurun_out/gather_ncu.log 2>&1
python scripts/profile_jk.py > gpurun_out/jk_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:jk_rows -s 1 -c 1 -o gpurun_out/jk_full_r1g python scripts/profile_jk.py > gpurun_out/jk_ncu.log 2>&1
cat gpurun_out/jk_plain.log gpurun_out/k3_plain.log; tail -n 4 gpurun_out/sweep_g.log
