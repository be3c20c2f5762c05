#!/bin/bash
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
for n in 1 2 3 4 2 3 4; do
  HB_STREAMS=$n timeout 500 python bench.py --no-cpu-baseline > $O/${TAG}_s$n.json 2> $O/${TAG}_s$n.err; echo "streams=$n rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("$O/${TAG}_s$n.json").read().strip().splitlines()[-1])
    print("  value", d["value"], "e2e", d["e2e"]["value"])
except Exception as ex:
    print("  no line", ex)
PY
done
