#!/bin/bash
# GPU-box pass: decoder tests, timing, kernel durations under ncu, SASS evidence.  usage: bash tools/gpu_mlp_prof.sh TAG
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
timeout 200 python -m pytest tests/test_decoders_gpu.py -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
timeout 120 python tools/time_decoders.py > $O/${TAG}_time_decoders.json 2>/dev/null; echo "time rc=$?"; cat $O/${TAG}_time_decoders.json
timeout 300 ncu --clock-control none -k regex:'mlp_' -c 40 --metrics gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_tc.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active \
  --csv --log-file $O/${TAG}_mlp_ncu.csv python tools/time_decoders.py > /dev/null 2>&1; echo "ncu rc=$?"
grep -c mlp_ $O/${TAG}_mlp_ncu.csv
