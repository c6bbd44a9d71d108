set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest exit $?"; tail -n 5 gpurun_out/pytest_gpu_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/smoke_final.log
timeout 200 python scripts/jk_sweep.py --out gpurun_out/jk_sweep_final.json > gpurun_out/jk_sweep_final.log 2>&1; echo "jk_sweep exit $?"; cut -c1-200 gpurun_out/jk_sweep_final.log
timeout 900 python bench.py > gpurun_out/bench_r1q.json 2> gpurun_out/bench_r1q.err; echo "bench exit $?"; python -c "
import json;d=json.loads(open('gpurun_out/bench_r1q.json').read().strip().splitlines()[-1]);print(d['value'],d['e2e']['value'],d['roofline']['frac'],d['consumer'],d['clocks'])"
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref_r1q.json 2> gpurun_out/bench_ref_r1q.err; echo "ref exit $?"; cut -c1-400 gpurun_out/bench_ref_r1q.json
