"""One profiled frame of the bench workload (GPU box, under ncu --profile-from-start off).
usage: python tools/prof_frame.py [workload] [binning_mode]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench
from hugs_b200 import _native


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "human_scene_1080p"
    mode = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    _native.set_binning_mode(mode)
    rep = bench.Replica(bench.make_workload(workload, 0, 1, dev), dev, bench.our_impl(), sh_sink=True)
    rep.n_streams = 1
    rep.use_graph = False   # one eager frame on one stream: every kernel is a plain launch for ncu
    for _ in range(3):
        rep.step_resident()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    rep.step_resident()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    nd = rep.last_nodes[0]
    print("profiled one frame: P=%d R=%s mode=%d" % (rep.P, nd if isinstance(nd, int) or nd is None else nd.num_rendered,
                                                     mode))  # views=1


if __name__ == "__main__":
    main()
