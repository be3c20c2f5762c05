import os, sys, time, gc
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
module, lbs_impl = bench.our_impl()
rep = bench.Replica(bench.make_workload("human_scene_1080p", 0, 1, dev), dev, module, lbs_impl)
for _ in range(10): rep.step_resident()
torch.cuda.synchronize()
for label in ("gc on", "gc off", "gc on", "gc off"):
    if label == "gc off": gc.collect(); gc.disable()
    else: gc.enable()
    cpu = []
    torch.cuda.synchronize(); t00 = time.perf_counter()
    for _ in range(200):
        t0 = time.perf_counter(); rep.step_resident(); cpu.append(time.perf_counter() - t0)
    torch.cuda.synchronize(); t_all = time.perf_counter() - t00
    big = [(i, round(c * 1e3, 2)) for i, c in enumerate(cpu) if c > 4e-3]
    print(label, "wall %.3f ms/step" % (t_all / 200 * 1e3), "steps > 4 ms:", big, "gc counts", gc.get_count())
