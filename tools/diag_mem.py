import gc, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
V = int(sys.argv[1]) if len(sys.argv) > 1 else 4
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
module, lbs_impl = bench.our_impl()
rep = bench.Replica(bench.make_workload("human_scene_1080p", 0, 1, dev, views=V), dev, module, lbs_impl)
for i in range(40):
    rep.step_resident()
    if i % 5 == 0:
        torch.cuda.synchronize()
        st = torch.cuda.memory_stats()
        print(i, "allocated MiB", torch.cuda.memory_allocated() >> 20, "max_alloc", torch.cuda.max_memory_allocated() >> 20,
              "reserved", torch.cuda.memory_reserved() >> 20, "segments", st["segment.all.current"],
              "inactive_split MiB", st["inactive_split_bytes.all.current"] >> 20, "R", [int(n.num_rendered) for n in rep.last_nodes])
