#!/bin/bash
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
timeout 1700 python -m pytest tests -x -q -m gpu > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
run() { local name=$1; shift
  env "$@" timeout 500 python bench.py --no-cpu-baseline $ARGS > $O/${TAG}_$name.json 2> $O/${TAG}_$name.err; echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("$O/${TAG}_$name.json").read().strip().splitlines()[-1])
    print("  value", d["value"], "e2e", d["e2e"]["value"])
except Exception as ex:
    print("  no line", ex)
PY
  grep -m2 "Error" $O/${TAG}_$name.err | cut -c1-200
}
ARGS="--workload human_scene_separate_1080p" run sep1 HB_DEFER=1
ARGS="--workload human_scene_separate_1080p" run sep0 HB_DEFER=0
ARGS="--steps 200 --warmup 3" run long1 HB_DEFER=1
ARGS="--steps 200 --warmup 3" run long2 HB_DEFER=1
