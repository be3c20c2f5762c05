#!/bin/bash
# GPU-box pass: one bench line per BASELINE config, both arms.   usage: bash tools/gpu_configs.sh TAG
TAG=${1:-rXX}
O=gpurun_out
mkdir -p $O
run() {  # workload steps warmup ref_steps [extra args]
  local wl=$1 st=$2 wu=$3 rst=$4; shift 4
  timeout 900 python bench.py --workload $wl --steps $st --warmup $wu --cpu-baseline-seconds 10 "$@" \
     > $O/${TAG}_${wl}.json 2> $O/${TAG}_${wl}.err; echo "$wl ours rc=$?"
  timeout 900 python bench.py --impl reference --workload $wl --steps $rst --warmup 3 "$@" \
     > $O/${TAG}_${wl}_ref.json 2> $O/${TAG}_${wl}_ref.err; echo "$wl ref rc=$?"
  python - <<PY
import json
for suffix in ("", "_ref"):
    try:
        d = json.load(open("$O/${TAG}_${wl}%s.json" % suffix))
        print("  %-10s value %9.2f e2e %9.2f %s" % (suffix or "ours", d["value"], d["e2e"]["value"],
              {k: v["ms"] for k, v in d.get("stages", {}).items()}))
    except Exception as ex:
        print("  %s: no line (%r)" % (suffix or "ours", ex))
PY
}
run human_scene_1080p 100 10 12
run human_scene_separate_1080p 50 5 8
run human_only_1080p 100 10 20
run scene_only_3M_forward 50 5 10
run pose_animation_300 8 3 3
run human_only_256 100 10 20
ls -la $O | tail -20
