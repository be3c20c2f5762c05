#!/bin/bash
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'mlp_chain' -c 2 -f -o $O/${TAG}_mlp python tools/time_decoders.py > $O/${TAG}_mlp_full.log 2>&1; echo "rc=$?"
ncu -i $O/${TAG}_mlp.ncu-rep --page details > $O/${TAG}_mlp_details.txt 2>/dev/null
ncu -i $O/${TAG}_mlp.ncu-rep --page source --csv > $O/${TAG}_mlp_source.csv 2>/dev/null
ls -la $O | grep ${TAG}
