set -x
mkdir -p gpurun_out
(timeout 150 ./scripts/micro/k3_i8_proto_r6 20304 2716; echo "exit $?"; timeout 100 ./scripts/micro/k3_i8_proto 11936 2716; echo "exit $?"; timeout 100 ./scripts/micro/k3_i8_proto 5050 8000; echo "exit $?") > gpurun_out/k3_i8_proto3.txt 2>&1
cat gpurun_out/k3_i8_proto3.txt
