"""Host-side view of a bench step (GPU box): device time vs enqueue time, per-step outliers."""
import gc, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench

V = int(sys.argv[1]) if len(sys.argv) > 1 else 4
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
module, lbs_impl = bench.our_impl()
rep = bench.Replica(bench.make_workload("human_scene_1080p", 0, 1, dev, views=V), dev, module, lbs_impl)
for _ in range(5): rep.step_resident()
torch.cuda.synchronize(); gc.collect(); gc.disable()
for trial in range(3):
    cpu = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); t00 = time.perf_counter(); e0.record()
    for _ in range(25):
        t0 = time.perf_counter(); rep.step_resident(); cpu.append(time.perf_counter() - t0)
    e1.record(); torch.cuda.synchronize()
    n = 25 * V
    print("trial %d V=%d: gpu %.3f ms/frame; host step times ms: min %.2f med %.2f max %.2f" % (
        trial, V, e0.elapsed_time(e1) / n, min(cpu) * 1e3, sorted(cpu)[12] * 1e3, max(cpu) * 1e3))
st = torch.cuda.memory_stats()
print("reserved MiB", torch.cuda.memory_reserved() >> 20, "alloc retries", st["num_alloc_retries"],
      "cudaMalloc calls", st.get("segment.all.allocated", -1), "frees", st.get("segment.all.freed", -1))
