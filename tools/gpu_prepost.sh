#!/bin/bash
# GPU-box pass for the N1/N2 rows: their tests + timing.   usage: bash tools/gpu_prepost.sh TAG
TAG=${1:-rXX}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_triplane_gpu.py tests/test_renderer_gpu.py -m gpu -q > $O/${TAG}_prepost_pytest.log 2>&1; echo "pytest rc=$?"
tail -25 $O/${TAG}_prepost_pytest.log
timeout 600 python tools/time_pre_post.py > $O/${TAG}_prepost_timing.json 2> $O/${TAG}_prepost_timing.err; echo "timing rc=$?"
cat $O/${TAG}_prepost_timing.json; tail -5 $O/${TAG}_prepost_timing.err
