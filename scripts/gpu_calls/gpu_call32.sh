set -x
mkdir -p gpurun_out
timeout 300 python scripts/micro/ozaki_probe.py 11936 2720 > gpurun_out/ozaki_probe.txt 2>&1
timeout 300 python scripts/micro/ozaki_probe.py 20304 2720 >> gpurun_out/ozaki_probe.txt 2>&1
cat gpurun_out/ozaki_probe.txt
