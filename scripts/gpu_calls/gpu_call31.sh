set -x
mkdir -p gpurun_out
timeout 200 python scripts/micro/int8_gemm_rate.py > gpurun_out/int8_gemm_rate.txt 2>&1; cat gpurun_out/int8_gemm_rate.txt
