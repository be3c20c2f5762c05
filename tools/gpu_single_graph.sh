#!/bin/bash
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
run() { local name=$1; shift
  env "$@" timeout 500 python bench.py --no-cpu-baseline $ARGS > $O/${TAG}_$name.json 2> $O/${TAG}_$name.err; echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("$O/${TAG}_$name.json").read().strip().splitlines()[-1])
    print("  value", d["value"], "e2e", d["e2e"]["value"], "h2d", d["e2e"]["h2d_bytes_per_step"])
except Exception as ex:
    print("  no line", ex)
PY
  grep -m2 "Error" $O/${TAG}_$name.err | cut -c1-200
}
timeout 300 python -m pytest tests/test_graph_gpu.py -m gpu -q 2>&1 | tail -3
ARGS="--workload human_only_1080p" run ho_g1 HB_GRAPH=1
ARGS="--workload human_only_1080p" run ho_g0 HB_GRAPH=0
ARGS="--workload human_only_1080p --steps 150 --warmup 3" run ho_g1_long HB_GRAPH=1
ARGS="--workload human_only_1080p --steps 150 --warmup 3" run ho_g1_long2 HB_GRAPH=1
ARGS="--workload human_only_256" run h256_g1 HB_GRAPH=1
ARGS="--workload human_only_256" run h256_g0 HB_GRAPH=0
ARGS="--workload human_only_256 --steps 150 --warmup 3" run h256_g1_long HB_GRAPH=1
ARGS="--workload human_only_256 --steps 150 --warmup 3" run h256_g1_long2 HB_GRAPH=1
ARGS="--workload human_only_256 --steps 150 --warmup 3" run h256_g1_long3 HB_GRAPH=1
