#!/bin/bash
# 1 -> 8 GPU scaling of the bench on one box (run under gpurun --gpus 8).
TAG=${1:-rXX}
O=gpurun_out
mkdir -p $O
nvidia-smi topo -m > $O/${TAG}_topo.txt 2>&1
for N in ${NS:-1 2 4 8}; do
  if [ $N = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 30 --warmup 5 --no-cpu-baseline > $O/${TAG}_scale_$N.json 2> $O/${TAG}_scale_$N.err
  else
    NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $N --steps 30 --warmup 5 > $O/${TAG}_scale_$N.out 2> $O/${TAG}_scale_$N.err
    grep '^{' $O/${TAG}_scale_$N.out > $O/${TAG}_scale_$N.json
    grep -h -m3 -i "nvls" $O/${TAG}_scale_$N.out $O/${TAG}_scale_$N.err | cut -c1-200
  fi
  echo "N=$N rc=$?"; python -c "
import json;d=json.load(open('$O/${TAG}_scale_$N.json'));print($N,d['value'],d['e2e']['value'],d['ms_per_step'])"
done
python - <<PY
import json
for n in "${NS:-1 2 4 8}".split():
    d=json.load(open("$O/${TAG}_scale_%s.json" % n))
    print(n, "dp:", d.get("dp"))
PY
