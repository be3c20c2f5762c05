#!/bin/bash
# ncu passes on the GPU box: launch list of a short bench run + one --set full frame of our kernels.
# usage: bash tools/gpu_ncu.sh TAG   (outputs under gpurun_out/, kept below the 64 MiB merge limit)
TAG=${1:-rXX}
O=gpurun_out
mkdir -p $O
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:'render_|preprocess|onesweep|duplicate|scan_kernel|tile_ranges|lbs_|histogram|offsets' \
  -f -o $O/${TAG}_frame python tools/prof_frame.py > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -2 $O/${TAG}_ncu_full.log
ncu -i $O/${TAG}_frame.ncu-rep --page raw --csv > $O/${TAG}_frame_raw.csv 2>/dev/null
ncu -i $O/${TAG}_frame.ncu-rep --page details > $O/${TAG}_frame_details.txt 2>/dev/null
SZ=$(stat -c %s $O/${TAG}_frame.ncu-rep); echo "rep bytes $SZ"
if [ "$SZ" -gt 45000000 ]; then rm -f $O/${TAG}_frame.ncu-rep; echo "rep dropped (too large)"; fi
du -sh $O
