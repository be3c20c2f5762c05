"""Bisect of the CUDA-graph replay fault (two streams inside one captured step, single-model path).

    python tools/graph_bisect.py VARIANT [steps]      one variant in this process
    python tools/graph_bisect.py all [steps]          every variant, each in its own process (twice)

A variant builds bench.Replica for a 256x256 workload, captures the resident step (hugs_b200.dp.GraphedStep;
HB_PIN_ACC=0 switches dp_step's pinning of the AccumulateGrad nodes off, which is what made it fail), replays it `steps` times with a device synchronize after every replay
(so the failing step is known) and compares the parameter gradients of every replay with the eager step's.
Result of the round-2 run: profiles/r02_aw_graph_bisect.txt."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

VARIANTS = {
    # name: (workload overrides, env, views)
    "base": (dict(), dict(), 8),
    "one_stream": (dict(), dict(HB_STREAMS="1"), 8),
    "two_views": (dict(), dict(), 2),
    "fwd_only": (dict(mode="forward"), dict(), 8),
    "scene_only": (dict(n_human=0, n_scene=110210), dict(), 8),
    "two_seg": (dict(n_human=60000, n_scene=50210), dict(), 8),
    "eager": (dict(), dict(BISECT_EAGER="1"), 8),
}


def run(variant, steps):
    over, env, views = VARIANTS[variant]
    os.environ.update(env)
    os.environ["HB_GRAPH"] = "0"   # the tool does its own capture
    import torch
    import bench
    from hugs_b200 import dp
    wl_name = "bisect_" + variant
    bench.WORKLOADS[wl_name] = dict(bench.WORKLOADS["human_only_256"], **over)
    device = torch.device("cuda", 0)
    torch.cuda.set_device(device)
    wl = bench.make_workload(wl_name, 0, 1, device, views=views)
    rep = bench.Replica(wl, device, bench.our_impl(), sh_sink=True)
    rep.step_resident()
    torch.cuda.synchronize()

    def grads():
        if rep.buf is None:
            return None
        if not rep.buf.gather:
            return rep.buf.flat.clone()
        return torch.cat([g.detach().reshape(-1).clone() for g in rep.buf.grads().values() if g is not None])

    ref = grads()
    print("%s: streams=%d two_seg=%s views=%d" % (variant, rep.n_streams, rep.two_seg, rep.views), flush=True)
    graphed, adopted = None, None
    worst = 0.0
    for k in range(steps):
        if env.get("BISECT_EAGER") == "1":
            rep.step_resident()
        else:
            if graphed is None:
                graphed = dp.GraphedStep(lambda: rep._step_resident_eager(exchange=False))
            graphed.replay()
            if rep.buf is not None and rep.buf.gather:   # the graph writes the tensors autograd adopted in capture
                if adopted is None:
                    adopted = {n: p.grad for n, p in rep.buf.params.items() if n not in rep.buf.sunk}
                for n, g in adopted.items():
                    rep.buf.params[n].grad = g
        torch.cuda.synchronize()
        if ref is not None:
            err = float((grads() - ref).abs().max() / ref.abs().max().clamp_min(1e-30))
            worst = max(worst, err)
            if err > 1e-3:
                print("  step %d: gradient differs from the eager step by %.3e (relative to max)" % (k, err), flush=True)
        if k % 20 == 0:
            print("  step %d ok" % k, flush=True)
    print("%s: done, %d steps, worst gradient difference %.3e" % (variant, steps, worst), flush=True)


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    if what != "all":
        return run(what, steps)
    for name in VARIANTS:
        for attempt in range(2):
            p = subprocess.run([sys.executable, os.path.abspath(__file__), name, str(steps)], capture_output=True,
                               text=True, timeout=300)
            tail = [ln for ln in (p.stdout + p.stderr).splitlines() if ln.strip()][-4:]
            print("== %s #%d rc=%d" % (name, attempt, p.returncode))
            for ln in tail:
                print("   " + ln[:200])
            sys.stdout.flush()


if __name__ == "__main__":
    main()
