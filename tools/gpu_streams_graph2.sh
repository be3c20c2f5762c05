#!/bin/bash
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
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
ARGS="" run s8 HB_STREAMS=8
ARGS="" run s6 HB_STREAMS=6
ARGS="" run s4_eager HB_STREAMS=4 HB_GRAPH=0
ARGS="--steps 150 --warmup 3" run s4_long HB_STREAMS=4
ARGS="--steps 150 --warmup 3" run s4_long2 HB_STREAMS=4
ARGS="--steps 100 --warmup 3 --workload human_scene_separate_1080p" run s4_sep HB_STREAMS=4
ARGS="--workload human_scene_separate_1080p" run s2_sep HB_STREAMS=2
