# Round 2 (1 GPU): consumer pipeline host timeline (e2e_consumer varied 1.84 - 2.55 s between boxes), the new int8
# guaranteed-tolerance test, bench N = 1 with the int8 extra.
set -x
mkdir -p gpurun_out
nproc; grep -m1 "model name" /proc/cpuinfo
timeout 300 python scripts/consumer_timeline.py > gpurun_out/r2c10_consumer_timeline.txt 2>&1; echo "consumer timeline exit $?"; cat gpurun_out/r2c10_consumer_timeline.txt | tail -n 9
OMP_NUM_THREADS=4 timeout 300 python scripts/consumer_timeline.py > gpurun_out/r2c10_consumer_timeline_omp4.txt 2>&1; echo "omp4 exit $?"; tail -n 9 gpurun_out/r2c10_consumer_timeline_omp4.txt
timeout 600 python -m pytest tests -m gpu -q -x -k "k3_int8 or seam" > gpurun_out/r2c10_pytest.log 2>&1; echo "pytest exit $?"; tail -n 4 gpurun_out/r2c10_pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/r2c10_bench.json 2> gpurun_out/r2c10_bench.err; echo "bench exit $?"; python -c "
import json; d=json.loads(open('gpurun_out/r2c10_bench.json').read().strip().splitlines()[-1]); print('value',d['value'],'e2e',d['e2e']['value'],'consumer',d['e2e_consumer']['value'],'traffic',d['roofline']['traffic']); print(d['k3_int8_extra'])"; tail -n 3 gpurun_out/r2c10_bench.err
