#!/bin/bash
# N = 1 and N = 8 with a given --views-per-rank (run under gpurun --gpus 8).  usage: bash tools/gpu_scale_v.sh TAG VIEWS
TAG=${1:-rXX}; V=${2:-8}
O=gpurun_out
mkdir -p $O
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --views-per-rank $V > $O/${TAG}_v${V}_1.json 2> $O/${TAG}_v${V}_1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 \
  bench.py --gpus 8 --steps 20 --warmup 5 --views-per-rank $V > $O/${TAG}_v${V}_8.out 2> $O/${TAG}_v${V}_8.err
grep '^{' $O/${TAG}_v${V}_8.out > $O/${TAG}_v${V}_8.json
python - <<PY
import json
a=json.load(open("$O/${TAG}_v${V}_1.json")); b=json.load(open("$O/${TAG}_v${V}_8.json"))
print("views", $V, "N=1", a["value"], a["e2e"]["value"], a["ms_per_step"], "N=8", b["value"], b["e2e"]["value"], b["ms_per_step"], "ratio", round(b["value"]/a["value"],3), "e2e ratio", round(b["e2e"]["value"]/a["e2e"]["value"],3))
print(b.get("dp"))
PY
