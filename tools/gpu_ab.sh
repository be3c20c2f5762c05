#!/bin/bash
# A/B of environment switches on the headline workload.  usage: bash tools/gpu_ab.sh TAG "ENV1=.." "ENV2=.." ...
TAG=${1:-rXX}; shift
O=gpurun_out; mkdir -p $O
i=0
for envs in "$@"; do
  i=$((i+1))
  env $envs timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > $O/${TAG}_ab$i.json 2> $O/${TAG}_ab$i.err; echo "[$envs] rc=$?"
  grep -c "AccumulateGrad" $O/${TAG}_ab$i.err
  python -c "
import json; d=json.load(open('$O/${TAG}_ab$i.json')); print('  value %.2f e2e %.2f frame %.4f' % (d['value'], d['e2e']['value'], d['frame']['ms']))"
done
