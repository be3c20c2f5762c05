#!/bin/bash
# Round-end check on one GPU: smoke, GPU suite, both bench arms with default flags.  usage: bash tools/gpu_final.sh TAG
TAG=${1:-rXX}
O=gpurun_out
mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/${TAG}_smoke.log
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/${TAG}_pytest.log
timeout 900 python bench.py --impl reference --steps 10 --warmup 3 > $O/${TAG}_ref.json 2> $O/${TAG}_ref.err; echo "ref rc=$?"
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
python - <<PY
import json
r=json.load(open("$O/${TAG}_ref.json")); d=json.load(open("$O/${TAG}_bench.json"))
print("ref", r.get("value"), r.get("e2e",{}).get("value"), r.get("ms_per_step"))
print("ours", d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["kernel"], d["roofline"]["frac"], d["cpu_baseline"]["value"], d["gpu_launches"], d["clocks"])
PY
