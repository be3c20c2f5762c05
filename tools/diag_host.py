"""Where does a bench step spend host time?  (GPU box)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from hugs_b200 import _native

dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
module, lbs_impl = bench.our_impl()
rep = bench.Replica(bench.make_workload("human_scene_1080p", 0, 1, dev), dev, module, lbs_impl)
for _ in range(10): rep.step_resident()
torch.cuda.synchronize()
for trial in range(3):
    cpu = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); t00 = time.perf_counter(); e0.record()
    for _ in range(50):
        t0 = time.perf_counter(); rep.step_resident(); cpu.append(time.perf_counter() - t0)
    e1.record(); t_enq = time.perf_counter() - t00; torch.cuda.synchronize(); t_all = time.perf_counter() - t00
    cpu.sort()
    print("trial %d: gpu %.3f ms/step, host enqueue total %.3f ms/step (median step %.3f, max %.3f), wall %.3f ms/step" % (
        trial, e0.elapsed_time(e1) / 50, t_enq / 50 * 1e3, cpu[25] * 1e3, cpu[-1] * 1e3, t_all / 50 * 1e3))
print(torch.cuda.memory_stats()["num_alloc_retries"], torch.cuda.memory_reserved() / 2**20, "MiB reserved")
# split a step
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(20): rep.step_resident()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
