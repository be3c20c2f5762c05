#!/bin/bash
# 2-GPU check of the graph-replayed step (allreduce outside the graph).  usage: gpurun --gpus 2 -- bash tools/gpu_n2.sh TAG
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
for g in 1 0; do
  HB_GRAPH=$g timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus 2 --steps 30 --warmup 5 > $O/${TAG}_n2_g$g.out 2> $O/${TAG}_n2_g$g.err; echo "N=2 graph=$g rc=$?"
  grep '^{' $O/${TAG}_n2_g$g.out > $O/${TAG}_n2_g$g.json
  python - <<PY
import json
try:
    d = json.load(open("$O/${TAG}_n2_g$g.json"))
    print("  value %.1f e2e %.1f ms/step %.3f dp %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], (d.get("dp") or {}).get("exchange_ms_per_step_by_rank")))
except Exception as ex:
    print("  no line", ex)
PY
  grep -m3 "Error\|error" $O/${TAG}_n2_g$g.err | cut -c1-200
done
