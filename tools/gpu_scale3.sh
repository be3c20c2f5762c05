#!/bin/bash
# Weak + strong scaling records on one 8-GPU box (run under gpurun --gpus 8).  usage: bash tools/gpu_scale2.sh TAG
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
run() {  # name N extra-args...
  local name=$1 N=$2; shift 2
  if [ $N = 1 ]; then
    timeout 600 python bench.py --gpus 1 --no-cpu-baseline "$@" > $O/${TAG}_${name}.json 2> $O/${TAG}_${name}.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $N "$@" > $O/${TAG}_${name}.out 2> $O/${TAG}_${name}.err
    grep '^{' $O/${TAG}_${name}.out > $O/${TAG}_${name}.json
  fi
  python - <<PY
import json
try:
    d = json.load(open("$O/${TAG}_${name}.json"))
    print("$name N=$N value %.1f e2e %.1f ms/step %.3f dp %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"],
          (d.get("dp") or {}).get("exchange_ms_per_step_by_rank")))
except Exception as ex:
    print("$name: no line", ex)
PY
}
run weak_1 1 --steps 30 --warmup 5
run weak_8 8 --steps 30 --warmup 5
run strong8_8 8 --steps 60 --warmup 8 --scaling strong --global-batch 8
