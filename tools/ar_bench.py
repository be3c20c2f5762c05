"""Allreduce micro-benchmark for the DP gradient exchange (275 MB fp32), run under torchrun.
Prints per-setting time; NCCL_ALGO etc. are taken from the environment of the launch."""
import os
import sys
import torch
import torch.distributed as dist

def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    n = 274793920 // 4
    x = torch.ones(n, device=dev)
    res = {}
    def timeit(fn, iters=20, warm=5):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters
    res["nccl_all_reduce"] = timeit(lambda: dist.all_reduce(x))
    x.fill_(1.0)
    try:
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(n, dtype=torch.float32, device=dev)
        symm.rendezvous(t, dist.group.WORLD.group_name)
        t.fill_(1.0)
        res["symm_multimem_all_reduce"] = timeit(lambda: torch.ops.symm_mem.multimem_all_reduce_(t, "sum", dist.group.WORLD.group_name))
        t.fill_(1.0)
        res["symm_two_shot_all_reduce"] = timeit(lambda: torch.ops.symm_mem.two_shot_all_reduce_(t, "sum", dist.group.WORLD.group_name))
    except Exception as e:  # noqa: BLE001
        res["symm_error"] = repr(e)[:200]
    if dist.get_rank() == 0:
        print("ALGO=%s PROTO=%s :" % (os.environ.get("NCCL_ALGO"), os.environ.get("NCCL_PROTO")),
              {k: (round(v, 3) if isinstance(v, float) else v) for k, v in res.items()}, flush=True)
    dist.destroy_process_group()

if __name__ == "__main__":
    main()
