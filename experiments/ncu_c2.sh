#!/bin/bash
# ncu evidence for the C2 workload (one GPU): launch list of the training kernels + full captures of the HBM-bound kernels.
# Usage (under gpurun): bash experiments/ncu_c2.sh <tag>
set -u
TAG=${1:-r02}
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity-check --no-profile"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_bt_|k_grad|k_quantize|k_scales|k_absmax|k_peer' -c 700 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_bt_hist -s 14 -c 14 -o gpurun_out/${TAG}_hist $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k2_bin_rows -s 50 -c 2 -o gpurun_out/${TAG}_k2 $CMD > gpurun_out/${TAG}_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_bt_partition|k_bt_apply_tree|k_quantize' -s 30 -c 16 -o gpurun_out/${TAG}_other $CMD > gpurun_out/${TAG}_ncu4.log 2>&1
ls -la gpurun_out | grep ${TAG}
