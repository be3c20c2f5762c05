#!/bin/bash
# ncu --set full of render_backward for the in-tree library and a variant.  usage: bash tools/gpu_ncu_bwd_ab.sh TAG variant
TAG=${1:-rXX}; VAR=$2; O=gpurun_out; mkdir -p $O
export HB_GRAPH=0
for name in base $VAR; do
  if [ $name = base ]; then unset HUGS_B200_LIB; else export HUGS_B200_LIB=$PWD/ml-hugs_b200/lib/variants/libhugs_b200_$name.so; fi
  timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'render_backward' \
    -f -o $O/${TAG}_$name python tools/prof_frame.py > $O/${TAG}_$name.log 2>&1; echo "ncu $name rc=$?"
  ncu -i $O/${TAG}_$name.ncu-rep --page raw --csv > $O/${TAG}_${name}_raw.csv 2>/dev/null
  ls -la $O/${TAG}_$name.ncu-rep | awk '{print $5}'
done
