#!/bin/bash
# One --set full capture of the kernels matching REGEX in one frame (tools/prof_frame.py).
# usage: bash tools/gpu_ncu_one.sh TAG REGEX
TAG=${1:-rXX}; RX=${2:-duplicate}
O=gpurun_out
mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"$RX" \
  -f -o $O/${TAG}_one python tools/prof_frame.py > $O/${TAG}_one_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i $O/${TAG}_one.ncu-rep --page raw --csv > $O/${TAG}_one_raw.csv 2>/dev/null
ncu -i $O/${TAG}_one.ncu-rep --page details > $O/${TAG}_one_details.txt 2>/dev/null
ls -la $O/${TAG}_one.ncu-rep
