#!/bin/bash
# GPU-box pass: bench lines of a few workloads, our arm only.   usage: bash tools/gpu_bench_wl.sh TAG wl1 wl2 ...
TAG=${1:-rXX}; shift
O=gpurun_out
mkdir -p $O
for wl in "$@"; do
  timeout 900 python bench.py --workload $wl --steps 50 --warmup 5 --no-cpu-baseline > $O/${TAG}_${wl}.json 2> $O/${TAG}_${wl}.err; echo "$wl rc=$?"
  tail -3 $O/${TAG}_${wl}.err
  python - <<PY
import json
try:
    d = json.load(open("$O/${TAG}_${wl}.json"))
    print("  value %9.2f e2e %9.2f frame %.4f ms %s" % (d["value"], d["e2e"]["value"], d["frame"]["ms"], {k: v["ms"] for k, v in d.get("stages", {}).items()}))
except Exception as ex:
    print("  no line (%r)" % (ex,))
PY
done
