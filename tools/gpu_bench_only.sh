#!/bin/bash
# bench line only (no tests, no CPU baseline).   usage: bash tools/gpu_bench_only.sh TAG
TAG=${1:-rXX}
O=gpurun_out
mkdir -p $O
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench.json"))
print("value", d["value"], "e2e", d["e2e"]["value"], "frame ms", d["frame"]["ms"], {k:v["ms"] for k,v in d["stages"].items()})
PY
