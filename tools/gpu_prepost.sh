#!/bin/bash
# GPU-box pass for the N1/N2 rows: their tests, timing, and (NCU=1) one --set full capture of their kernels.
# usage: [NCU=1] bash tools/gpu_prepost.sh TAG
TAG=${1:-rXX}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_triplane_gpu.py tests/test_renderer_gpu.py -m gpu -q > $O/${TAG}_prepost_pytest.log 2>&1; echo "pytest rc=$?"
tail -8 $O/${TAG}_prepost_pytest.log
timeout 600 python tools/time_pre_post.py > $O/${TAG}_prepost_timing.json 2> $O/${TAG}_prepost_timing.err; echo "timing rc=$?"
cat $O/${TAG}_prepost_timing.json; tail -5 $O/${TAG}_prepost_timing.err
if [ -n "$NCU" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:'triplane_|densification' -c 6 \
    --launch-skip 40 -f -o $O/${TAG}_prepost python tools/time_pre_post.py > $O/${TAG}_prepost_ncu.log 2>&1; echo "ncu rc=$?"
  ncu -i $O/${TAG}_prepost.ncu-rep --page raw --csv > $O/${TAG}_prepost_raw.csv 2>/dev/null
  ncu -i $O/${TAG}_prepost.ncu-rep --page details > $O/${TAG}_prepost_details.txt 2>/dev/null
  ls -la $O/${TAG}_prepost.ncu-rep
fi
