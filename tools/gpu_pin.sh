#!/bin/bash
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
for pin in 1 0; do
  for v in base two_views scene_only; do
    for rep in 1 2; do
      HB_PIN_ACC=$pin HB_STREAMS=2 timeout 300 python tools/graph_bisect.py $v 80 > $O/${TAG}_pin${pin}_${v}_$rep.txt 2>&1; rc=$?
      echo "pin=$pin $v #$rep rc=$rc $(grep -c differs $O/${TAG}_pin${pin}_${v}_$rep.txt) bad steps; $(tail -1 $O/${TAG}_pin${pin}_${v}_$rep.txt | cut -c1-100)"
    done
  done
done
HB_PIN_ACC=1 HB_STREAMS=4 timeout 300 python tools/graph_bisect.py base 80 2>&1 | tail -1
bash tools/gpu_ab.sh ${TAG} "HB_PIN_ACC=1" "HB_PIN_ACC=0" "HB_PIN_ACC=1 HB_GRAPH=0" "HB_PIN_ACC=0 HB_GRAPH=0" 2>&1 | grep -v "^0$"
