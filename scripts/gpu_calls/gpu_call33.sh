set -x
mkdir -p gpurun_out
timeout 60 ./scripts/micro/umma_i8_probe > gpurun_out/umma_i8_probe.txt 2>&1; echo "exit $?" >> gpurun_out/umma_i8_probe.txt
cat gpurun_out/umma_i8_probe.txt
nvidia-smi --query-gpu=name,memory.used --format=csv
