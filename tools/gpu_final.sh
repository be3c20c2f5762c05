#!/bin/bash
# What the driver runs at round end, on one box: GPU tests, smoke(), both bench arms with default arguments.
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/${TAG}_smoke.log
timeout 900 python bench.py --impl reference > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "ref rc=$?"
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; tail -2 $O/${TAG}_bench.err
python - <<PY
import json
a=json.load(open("$O/${TAG}_bench.json")); r=json.load(open("$O/${TAG}_bench_ref.json"))
print("ours", a["value"], a["e2e"]["value"], "ref", r["value"], r["e2e"]["value"], "ratio", a["value"]/r["value"], a["e2e"]["value"]/r["e2e"]["value"])
print(a["roofline"]["kernel"], a["roofline"]["bound"], a["roofline"]["frac"], a["roofline"].get("issue_frac"), a["cpu_baseline"])
print({k:v["ms"] for k,v in a["stages"].items()})
PY
