#!/bin/bash
# A/B of the e2e plumbing (G buffers, one small-record upload per step).  usage: bash tools/gpu_e2e_ab.sh TAG
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
run() { local name=$1; shift
  env "$@" timeout 400 python bench.py --no-cpu-baseline > $O/${TAG}_$name.json 2> $O/${TAG}_$name.err; echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("$O/${TAG}_$name.json").read().strip().splitlines()[-1])
    print("  value", d["value"], "e2e", d["e2e"]["value"])
except Exception as ex:
    print("  no line", ex)
PY
}
run old HB_E2E_GBUF=2 HB_E2E_SMALL_ONCE=0
run new HB_E2E_GBUF=4 HB_E2E_SMALL_ONCE=1
run buf4 HB_E2E_GBUF=4 HB_E2E_SMALL_ONCE=0
run small HB_E2E_GBUF=2 HB_E2E_SMALL_ONCE=1
run old2 HB_E2E_GBUF=2 HB_E2E_SMALL_ONCE=0
run new2 HB_E2E_GBUF=4 HB_E2E_SMALL_ONCE=1
timeout 300 python -m pytest tests/test_bench_gpu.py -m gpu -q 2>&1 | tail -2
