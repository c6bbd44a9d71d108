# Round 2 (1 GPU), last seconds: K3 DRAM traffic with wider bands whose row panels fit the L2 persisting set-aside.
mkdir -p gpurun_out
IFS=";" read -ra CFGS <<< "${QEMB_CFGS:-5 24 72;5 28 79}"
for cfg in "${CFGS[@]}"; do IFS=" " read -r a b c <<< "$cfg"; set -- $a $b $c
  QEMB_K3_HINT=$1 QEMB_K3_GROUP=$2 QEMB_L2_PERSIST_MB=$3 timeout 38 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct -k regex:gram_kernel --clock-control none -s 1 -c 1 --csv \
    --log-file gpurun_out/r2c20_k3_$1_$2_$3.csv python scripts/profile_k3.py 25425 2716 1 > /dev/null 2>&1; echo "ncu hint $1 band $2 persist $3 exit $?"
  grep -E "dram__bytes|hit_rate|gpu__time" gpurun_out/r2c20_k3_$1_$2_$3.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' | sed "s/^/hint $1 band $2 persist $3: /" | tee -a gpurun_out/r2c20_k3_traffic.txt
done
