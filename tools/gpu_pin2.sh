#!/bin/bash
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
timeout 1700 python -m pytest tests -x -q -m gpu > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
run() { local name=$1; shift
  timeout 500 python bench.py --no-cpu-baseline "$@" > $O/${TAG}_$name.json 2> $O/${TAG}_$name.err; echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("$O/${TAG}_$name.json").read().strip().splitlines()[-1])
    print("  value", d["value"], "e2e", d["e2e"]["value"], d["impl"]["cuda_streams_per_rank"])
except Exception as ex:
    print("  no line", ex)
PY
  grep -m2 "Error" $O/${TAG}_$name.err | cut -c1-200
}
run ho --workload human_only_1080p
run ho_long --workload human_only_1080p --steps 200 --warmup 3
run ho_long2 --workload human_only_1080p --steps 200 --warmup 3
run h256 --workload human_only_256
run h256_long --workload human_only_256 --steps 200 --warmup 3
run h256_long2 --workload human_only_256 --steps 200 --warmup 3
run sep --workload human_scene_separate_1080p
run head
