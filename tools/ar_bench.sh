for A in "" "NVLS" "Ring" "NVLSTree"; do
  NCCL_ALGO=$A timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tools/ar_bench.py 2>&1 | grep "ALGO=" 
done
