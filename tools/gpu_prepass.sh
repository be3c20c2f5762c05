#!/bin/bash
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_grad_sinks_gpu.py tests/test_graph_gpu.py tests/test_bench_gpu.py tests/test_parity_gpu.py -x -q -m gpu > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/${TAG}_pytest.log
bash tools/gpu_ab.sh $TAG "HB_PREPASS=1" "HB_PREPASS=0" "HB_PREPASS=1 HB_GRAPH=0" "HB_PREPASS=0 HB_GRAPH=0" 2>&1 | grep -v "^[01]$"
python - <<PY
import json
for i in (1,2):
    d=json.load(open("$O/${TAG}_ab%d.json"%i)); print({k: v["ms"] for k, v in d["stages"].items()})
PY
