set -x
mkdir -p gpurun_out
timeout 150 ./scripts/micro/k3_i8_proto 20304 2716 > gpurun_out/k3_i8_proto.txt 2>&1; echo "exit $?" >> gpurun_out/k3_i8_proto.txt
cat gpurun_out/k3_i8_proto.txt
nvidia-smi --query-gpu=name,memory.used --format=csv,noheader
