#!/bin/bash
# GPU-box pass: GPU test suite + one bench line.   usage: bash tools/gpu_quick.sh TAG [bench args]
TAG=${1:-rXX}; shift
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
tail -12 $O/${TAG}_pytest.log
timeout 900 python bench.py --steps 50 --warmup 5 "$@" > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
tail -3 $O/${TAG}_bench.err
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench.json"))
print("value", d["value"], "e2e", d["e2e"]["value"], "frame ms", d["frame"]["ms"], {k:(v["ms"],v["frac"]) for k,v in d["stages"].items()})
PY
