#!/bin/bash
# ncu full capture of selected kernels on a small forward; exports the raw page as CSV and per-instance
# source-page digests (tools/ncu_regions.py, tools/ncu_hot.py); the .ncu-rep stays on the box (size).
# usage: gpu_ncu_full.sh <kernel-regex> <count> "<instance ids for source digests>" [workload] [batch]
mkdir -p gpurun_out
K=${1:-tc_rows_kernel}; C=${2:-8}; IDS=${3:-0}; WL=${4:-C2}; B=${5:-256}
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$K" -c $C -o /tmp/cap -f python tools/one_forward.py $WL $B bf16 > gpurun_out/ncu_full.log 2>&1; echo full rc=$?
ncu -i /tmp/cap.ncu-rep --page raw --csv > gpurun_out/cap_raw.csv 2>/dev/null
for i in $IDS; do
  ncu -i /tmp/cap.ncu-rep --page source --csv --launch-skip $i --launch-count 1 > /tmp/src_$i.csv 2>/dev/null
  python tools/ncu_regions.py /tmp/src_$i.csv 0 24 150 > gpurun_out/regions_$i.txt 2>&1
  python tools/ncu_hot.py /tmp/src_$i.csv 0 > gpurun_out/hot_$i.txt 2>&1
done
ls -la gpurun_out/ | head -30
