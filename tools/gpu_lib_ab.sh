#!/bin/bash
# A/B of library variants (tools/build_variant.sh) on the headline workload.  usage: bash tools/gpu_lib_ab.sh TAG name1 name2 ...
# ("base" = the in-tree library)
TAG=${1:-rXX}; shift
O=gpurun_out; mkdir -p $O
for name in "$@"; do
  if [ $name = base ]; then unset HUGS_B200_LIB; else export HUGS_B200_LIB=$PWD/ml-hugs_b200/lib/variants/libhugs_b200_$name.so; fi
  timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > $O/${TAG}_$name.json 2> $O/${TAG}_$name.err; echo "[$name] rc=$?"
  python -c "
import json; d=json.load(open('$O/${TAG}_$name.json')); print('  value %.2f e2e %.2f' % (d['value'], d['e2e']['value']), {k: v['ms'] for k, v in d['stages'].items()})"
done
