#!/bin/bash
# GPU-box pass: the no-alloc test alone, the graph tests, graph on/off bench A/B, then the whole GPU suite.
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/${TAG}_noalloc_alone.log 2>&1; echo "no_alloc alone rc=$?"; tail -3 $O/${TAG}_noalloc_alone.log
timeout 600 python -m pytest tests/test_graph_gpu.py tests/test_bench_gpu.py -m gpu -q > $O/${TAG}_graph_tests.log 2>&1; echo "graph+bench tests rc=$?"; tail -5 $O/${TAG}_graph_tests.log
for g in 1 0 1 0; do
  HB_GRAPH=$g timeout 400 python bench.py --no-cpu-baseline > $O/${TAG}_bench_g$g.json 2> $O/${TAG}_bench_g$g.err; echo "bench graph=$g rc=$?"
  python - <<PY
import json
d=json.loads(open("$O/${TAG}_bench_g$g.json").read().strip().splitlines()[-1])
print("  value", d["value"], "e2e", d["e2e"]["value"], "launches", d.get("gpu_launches"))
PY
done
HB_GRAPH=1 timeout 600 python bench.py --steps 150 --warmup 3 --no-cpu-baseline --workload human_scene_separate_1080p > $O/${TAG}_sep_g1.json 2> $O/${TAG}_sep_g1.err; echo "separate graph stress rc=$?"
timeout 1700 python -m pytest tests -x -q -m gpu > $O/${TAG}_pytest_full.log 2>&1; echo "full suite rc=$?"; tail -4 $O/${TAG}_pytest_full.log
