#!/bin/bash
# GPU-box pass: selected GPU tests only.   usage: bash tools/gpu_tests_only.sh TAG <pytest args>
TAG=${1:-rXX}; shift
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest "$@" -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
tail -25 $O/${TAG}_pytest.log
