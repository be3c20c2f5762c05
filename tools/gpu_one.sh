#!/bin/bash
# GPU-box pass: a few tests under a short timeout (new kernels: a hang must not become a strike).
# usage: bash tools/gpu_one.sh TAG SECONDS <pytest args>
TAG=${1:-rXX}; T=${2:-120}; shift 2
O=gpurun_out; mkdir -p $O
timeout $T python -m pytest "$@" -m gpu -q -x > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
tail -40 $O/${TAG}_pytest.log
nvidia-smi --query-gpu=name,memory.used --format=csv,noheader
