#!/bin/bash
# ncu passes on the GPU box (after the plain runs have exited 0): launch list of a short bench run, one
# --set full frame, and the pipe / instruction counters of the two blend kernels.   usage: bash tools/gpu_prof2.sh TAG
TAG=${1:-rXX}
O=gpurun_out
mkdir -p $O
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_plain.json 2> $O/${TAG}_plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:'render_|preprocess|onesweep|duplicate|scan|tile_ranges|lbs_|histogram|offsets|l2_|layer_' \
  -f -o $O/${TAG}_frame python tools/prof_frame.py > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -2 $O/${TAG}_ncu_full.log
ncu -i $O/${TAG}_frame.ncu-rep --page raw --csv > $O/${TAG}_frame_raw.csv 2>/dev/null
timeout 900 ncu --clock-control none --profile-from-start off -k regex:'render_' \
  --metrics smsp__inst_executed.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_fmalite.sum,sm__inst_executed_pipe_uniform.sum,sm__inst_executed_pipe_cbu.sum,sm__inst_executed_pipe_adu.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio \
  --csv --log-file $O/${TAG}_blend_pipes.csv python tools/prof_frame.py > $O/${TAG}_ncu_pipes.log 2>&1; echo "ncu pipes rc=$?"
SZ=$(stat -c %s $O/${TAG}_frame.ncu-rep); echo "rep bytes $SZ"
if [ "$SZ" -gt 30000000 ]; then rm -f $O/${TAG}_frame.ncu-rep; echo "rep dropped (too large)"; fi
du -sh $O
