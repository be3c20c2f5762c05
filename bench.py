#!/usr/bin/env python3
"""bench.py -- fwd+bwd frames/s of the HUGS splat hot path on N B200s (BASELINE.json metric).

A "step" on each rank is --views-per-rank (default 8) frames of the human+scene workload (BASELINE.json configs[2]: 110 210
LBS-deformed human Gaussians + 1 000 000 scene Gaussians, SH degree 3, 1920x1080):
    LBS forward -> rasterizer forward -> loss=(color*G).sum() -> rasterizer backward -> LBS backward
per frame, gradients accumulated over the rank's frames, followed, for N > 1, by one NCCL
allreduce(sum) of the parameter gradients (view-batch data parallelism; a single view is never
split).  The step's view batch is N x views-per-rank views, so the per-rank work is fixed
("scaling": "weak") and frames/s = N x views-per-rank x steps / time.

  value    frames/s with every input already resident in HBM, CUDA-event timed, max over ranks
  e2e      the same step through the public API with the per-frame inputs (camera, SMPL joint
           transforms, the image-space loss gradient that stands in for the training target)
           copied from pinned host memory and the loss read back, inside the timed region (the
           25 MB target is uploaded on a copy stream under the forward pass; every frame's loss is
           copied to a pinned slot and read by the host four frames later, the pending ones before
           the closing event); the views of a step alternate between two CUDA streams
  roofline the dominant kernel's algorithmic bytes / its CUDA-event time / measured HBM peak
  stages   the same for every stage (per-stage CUDA events recorded inside the library on the
           launching stream during a second pass of K steps)
  cpu_baseline   the CPU oracle port timed on this box's host cores on a bounded sample (N=1 only)

`--impl reference` runs the UNMODIFIED reference (oracle/_ref: the reference's CUDA rasterizer built
from /root/reference + its PyTorch LBS chain) through its own public API on the same workload; if
that build is absent it times the CPU oracle port instead.  Nothing on our arm touches oracle/.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "ml-hugs_b200"), os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

METRIC = "fwd+bwd frames/s, 1.1M Gaussians @1080p"
UNIT = "frames/s"
# mode: "train" = LBS + rasterizer forward + loss + backward per view, gradients accumulated, allreduce per step
#       "train_separate" = the same with the trainer's separate human image (gs_renderer.py:68-82, two losses)
#       "forward" = rendering only (no_grad), the image is the result
#       "animation" = `frames` pose frames per step, split over the ranks: batched LBS + render, no collective
def _w(n_human, n_scene, W=1920, H=1080, focal=1400.0, mode="train", **kw):
    return dict(n_human=n_human, n_scene=n_scene, W=W, H=H, focal=focal, mode=mode, **kw)


WORKLOADS = {
    "human_scene_1080p": _w(110210, 1_000_000),                              # BASELINE.json configs[2] (the metric's)
    "human_only_1080p": _w(110210, 0),                                       # configs[1]
    "human_scene_separate_1080p": _w(110210, 1_000_000, mode="train_separate"),   # configs[2] as the trainer runs it
    "scene_only_3M_forward": _w(0, 3_000_000, mode="forward"),               # configs[3]
    "pose_animation_300": _w(110210, 0, mode="animation", frames=300),       # configs[4]
    "human_only_256": _w(110210, 0, 256, 256, 187.0, cpu="torch"),           # configs[0]: the CPU-runnable case
    "smoke_256": _w(4096, 20000, 256, 256, 187.0),
    "smoke_human_256": _w(4096, 0, 256, 256, 187.0),
    "smoke_separate_256": _w(4096, 20000, 256, 256, 187.0, mode="train_separate"),
    "smoke_forward_256": _w(0, 30000, 256, 256, 187.0, mode="forward"),
    "smoke_animation_256": _w(4096, 0, 256, 256, 187.0, mode="animation", frames=12),
}
SEED = 1002


def load_synth():
    """hugs_b200/synth.py loaded by path: pure torch, and the reference arm must not import the product
    package (importing hugs_b200 loads libhugs_b200.so)."""
    import importlib.util
    if "hb_synth" not in sys.modules:
        spec = importlib.util.spec_from_file_location(
            "hb_synth", os.path.join(ROOT, "ml-hugs_b200", "hugs_b200", "synth.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules["hb_synth"] = mod
        spec.loader.exec_module(mod)
    return sys.modules["hb_synth"]


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="human_scene_1080p", choices=sorted(WORKLOADS))
    ap.add_argument("--views-per-rank", type=int, default=8,
                    help="views each rank renders per step before the gradient allreduce (view batch = this x N)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --views-per-rank views on every rank; strong: --global-batch views split over the ranks")
    ap.add_argument("--global-batch", type=int, default=0, help="view batch of a step under --scaling strong")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-baseline-seconds", type=float, default=20.0)
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            c = [x.strip() for x in r.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0]))
                mx.append(float(c[1]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# workload
# ------------------------------------------------------------------------------------------------
def make_workload(name, rank, world, device, views=1):
    """Seeded synthetic tensors of the workload.  Parameters are identical on every
    rank (replicas); the step's view batch of world x views views on the orbit is split across the
    ranks, and every view has its own camera, SMPL joint transforms and image-space loss gradient.
    animation: the step's `frames` pose frames are split round-robin instead."""
    synth = load_synth()

    w = WORKLOADS[name]
    n_human, n_scene, W, H, focal, mode = w["n_human"], w["n_scene"], w["W"], w["H"], w["focal"], w["mode"]
    if mode == "animation":
        n_views = w["frames"]
        mine = list(range(rank, n_views, max(world, 1)))
    else:
        n_views = max(world, 1) * views
        # round-robin split of the step's view batch (hugs_b200.dp.shard_views): every rank gets the same
        # mix of orbit positions, so the ranks' pair counts -- and step times -- stay balanced
        mine = [v * max(world, 1) + rank for v in range(views)]
    cams = [synth.make_camera(W, H, focal, view=i, n_views=n_views, seed=SEED) for i in mine]
    parts = dict(cam=cams[0], cams=cams, W=W, H=H, n_human=n_human, n_scene=n_scene, mode=mode, view_ids=mine)
    if n_human > 0:
        parts["human"] = synth.make_human_gaussians(n_human, seed=SEED + 1)
        if mode == "animation":   # one smooth pose sequence, every rank takes its frames of it
            seq = synth.make_joint_transforms(n_views, SEED + 100)
            parts["As"] = [seq[i] for i in mine]
        else:
            parts["As"] = [synth.make_joint_transforms(1, SEED + 100 + i)[0] for i in mine]
        parts["A"] = parts["As"][0]
    if mode in ("train", "train_separate"):
        parts["Gs"] = [synth.make_dL_dcolor(W, H, SEED + i) for i in mine]
        parts["G"] = parts["Gs"][0]
    if mode == "train_separate":
        parts["Ghs"] = [synth.make_dL_dcolor(W, H, SEED + 5000 + i) for i in mine]
    if n_scene > 0:
        # scene frustum is laid out for the central camera so every rank holds the same replica
        cam0 = synth.make_camera(W, H, focal, seed=SEED)
        parts["scene"] = synth.make_scene_gaussians(n_scene, SEED + 2, cam0["tanfovx"], cam0["tanfovy"])
    return parts


class Replica:
    """Device-resident parameter replica + the per-frame inputs, and the frame step itself."""

    def __init__(self, wl, device, impl, sh_sink=False):
        self.device = device
        self.module = impl["module"]      # a diff_gaussian_rasterization-shaped module (ours or the reference)
        self.lbs_impl = impl["lbs"]       # callable(A, xyz, w, rotmat, scales, transl) -> (xyz', rotq, scales')
        self.lbs_frames = impl.get("lbs_frames")      # batched over pose frames (animation), or None
        self.render_separate = impl.get("separate")   # (settings, tensors, n_human, human_bg) -> (color, human_color, radii)
        self.render2 = impl.get("render2")   # our arm: two input segments, no torch.cat (hugs_b200/segments.py)
        self.mode = wl["mode"]
        self.training = self.mode in ("train", "train_separate")
        d = lambda t: t.to(device).contiguous()  # noqa: E731
        leaf = (lambda t: d(t).requires_grad_(True)) if self.training else d  # noqa: E731
        self.W, self.H = wl["W"], wl["H"]
        self.n_human, self.n_scene = wl["n_human"], wl["n_scene"]
        opac, shs, n_small = [], [], 0
        self.h_xyz = self.h_rotmat = self.h_scales = self.h_weights = self.transl = None
        if self.n_human > 0:
            h = wl["human"]
            self.h_xyz, self.h_rotmat = leaf(h["xyz"]), leaf(h["rotmat"])
            self.h_scales, self.h_weights = leaf(h["scales"]), leaf(h["lbs_weights"])
            self.transl = d(h["transl"][0])
            opac.append(h["opacity"])
            shs.append(h["shs"])
            n_small += sum(t.numel() for t in (h["xyz"], h["rotmat"], h["scales"], h["lbs_weights"]))
        self.s_means = self.s_scales = self.s_rots = None
        if self.n_scene > 0:
            sc = wl["scene"]
            self.s_means, self.s_scales, self.s_rots = leaf(sc["means3D"]), leaf(sc["scales"]), leaf(sc["rotations"])
            opac.append(sc["opacity"])
            shs.append(sc["shs"])
            n_small += sum(sc[k].numel() for k in ("means3D", "scales", "rotations"))
        # attributes that do not pass through LBS are kept concatenated (human first,
        # hugs/renderer/gs_renderer.py:33-37) -- except on our training arm, which hands the rasterizer the two
        # models as two segments
        self.P = sum(t.shape[0] for t in opac)
        self.buf = None
        # human + scene: the rasterizer gets the two models as two segments (no torch.cat) with gradient sinks on
        # every direct input (no per-view AccumulateGrad kernels); HB_TWO_SEG=0 switches back to the
        # concatenated call for comparison (measured: 751 vs 766 frames/s, 706 vs 742 end to end)
        # (a single model -- human only -- has nothing to concatenate: the plain rasterizer call is less host work)
        self.two_seg = (self.render2 is not None and self.training and self.n_human > 0 and self.n_scene > 0
                        and os.environ.get("HB_TWO_SEG", "1") == "1")
        if self.two_seg:
            self._init_two_segments(wl, device, leaf)
        else:
            self.opacity = leaf(torch.cat(opac))
            self.shs = leaf(torch.cat(shs))
            if self.render2 is not None and self.training:
                self._init_flat(device)
        if not hasattr(self, "sh_sink"):
            self.sh_sink = None
        self.comm_events = None  # [(start, end)] of the allreduce, when the staged pass asks for it
        # the device-resident step of the two-segment path (every gradient lands in the flat buffer that exists
        # before the capture) is replayed as ONE CUDA graph (hugs_b200.dp.GraphedStep); HB_GRAPH=0 runs it eagerly
        self.use_graph = (self.render2 is not None and self.training and self.buf is not None
                          and os.environ.get("HB_GRAPH", "1") == "1")
        self.graphed = None
        self._graph_grads = None
        # views of a step alternate between `n_streams` CUDA streams: the issue-bound blend kernels of one
        # view overlap the latency- and HBM-bound binning / per-Gaussian kernels of the next
        # (measured under graph replay, frames/s resident / end to end: 1 stream 756 / 734, 2: 807 / 797,
        # 3: 817 / 810, 4: 822 / 817, 6: 819 / 813, 8: 820 / 819; eager 4 streams 773 / 744)
        self.n_streams = max(1, int(os.environ.get("HB_STREAMS", "4"))) if sh_sink else 1
        # HB_PRIO=1: the views' streams are high-priority streams and the blend kernels are launched at the lowest
        # priority (hugs_b200_set_blend_priority): the other views' short kernels run in the CTA slots a blend
        # kernel frees instead of waiting behind its 8 160-CTA grid
        self.prio = self.render2 is not None and self.n_streams > 1 and os.environ.get("HB_PRIO", "1") == "1"
        if self.n_streams > 1 and self.render2 is not None:
            from hugs_b200 import dp
            self.streams = dp.view_streams(self.n_streams, device, prioritise=self.prio)
        else:
            self.streams = [torch.cuda.Stream(device) for _ in range(self.n_streams)] if self.n_streams > 1 else []
        self.means2D = torch.zeros(self.P, 3, device=device, requires_grad=self.training and not self.two_seg)
        # per-frame inputs: pinned host copies (for e2e) and resident device copies (for value)
        self.tanfovx, self.tanfovy = wl["cam"]["tanfovx"], wl["cam"]["tanfovy"]
        self.hosts, self.devs = [], []
        for i, cam in enumerate(wl["cams"]):
            host = {k: cam[k].clone().pin_memory() for k in ("viewmatrix", "projmatrix", "campos", "bg")}
            if self.n_human > 0:
                host["A"] = wl["As"][i].clone().pin_memory()
            if self.training:
                host["G"] = wl["Gs"][i].clone().pin_memory()
            if self.mode == "train_separate":
                host["Gh"] = wl["Ghs"][i].clone().pin_memory()
            self.hosts.append(host)
            self.devs.append({k: v.to(device) for k, v in host.items()})
        self.human_bg = torch.tensor([1.0, 1.0, 1.0], device=device)
        self.views = len(self.hosts)
        self.h2d_bytes = sum(v.numel() * v.element_size() for host in self.hosts for v in host.values())
        self.d2h_bytes = (4 + 4) * self.views if self.training else 3 * self.W * self.H * 4 * self.views
        self.last_nodes, self.last_radii_v = [None] * self.views, [None] * self.views
        self.last_node = self.last_radii = None
        # e2e plumbing (training): the image-space loss gradient (the stand-in for the training target, 25 MB)
        # is uploaded on a copy stream into one of two device buffers while the forward pass runs; the
        # loss of every step is copied to pinned memory and read one step later, so the host never
        # stalls the device (what a training loop's prefetcher / async logger does).
        self.copy_stream = torch.cuda.Stream(device)
        big = [k for k in ("G", "Gh") if k in self.hosts[0]]
        # (four buffers: with views on two streams, the upload for view k may start when view k-4 is done, a
        # whole view ahead of the forward pass that needs it; HB_E2E_GBUF=2 is round 1's double buffer)
        self.n_gbuf = max(2, int(os.environ.get("HB_E2E_GBUF", "4")))
        self.G_dev = [{k: torch.empty_like(self.devs[0][k]) for k in big} for _ in range(self.n_gbuf)]
        self.G_ready = [torch.cuda.Event() for _ in range(self.n_gbuf)]
        self.step_done = [None] * self.n_gbuf
        # the small per-view inputs (camera, joint transforms) of ALL of a step's views travel as ONE pinned
        # record at the start of the step: a per-view copy would queue behind the 25 MB uploads on the H2D engine
        self.prefetch_small = os.environ.get("HB_E2E_SMALL_ONCE", "1") == "1"
        self._small_all = None
        self.loss_host = [torch.zeros(1).pin_memory() for _ in range(4)]
        self.loss_ready = [None] * 4
        # e2e plumbing (rendering): every frame's image goes to one of four pinned host buffers
        self.img_host = [] if self.training else [torch.empty(3, self.H, self.W).pin_memory() for _ in range(4)]
        self.img_ready = [None] * 4
        self.images_read = 0
        self.e2e_step = 0
        self.losses = []
        self.e2e_blocks = None
        if self.use_graph:
            self._pack_inputs_into_blocks()

    def _pack_inputs_into_blocks(self):
        """Graph replay reads the step's inputs at fixed addresses: the per-view inputs of the whole step live in a
        few contiguous device blocks (the small ones in one [views, L] record, every field padded to 16 bytes),
        each with a pinned host block and a device staging block of the same shape (see _step_e2e_graphed)."""
        dev, V = self.device, self.views
        small_keys = [k for k in self.hosts[0] if k not in ("G", "Gh")]
        offs, off = {}, 0
        for k in small_keys:
            offs[k] = off
            off += (self.hosts[0][k].numel() + 3) // 4 * 4
        small_host = torch.zeros(V, off).pin_memory()
        for v in range(V):
            for k in small_keys:
                t = self.hosts[v][k]
                small_host[v, offs[k]:offs[k] + t.numel()] = t.reshape(-1)
        small_dev = small_host.to(dev)
        self.e2e_blocks = [(small_host, torch.empty_like(small_dev), small_dev)]
        for v in range(V):
            for k in small_keys:
                t = self.hosts[v][k]
                self.devs[v][k] = small_dev[v, offs[k]:offs[k] + t.numel()].view(t.shape)
        for k in ("G", "Gh"):
            if k in self.hosts[0]:
                host_all = torch.stack([h[k] for h in self.hosts]).pin_memory()
                dev_all = host_all.to(dev)
                self.e2e_blocks.append((host_all, torch.empty_like(dev_all), dev_all))
                for v in range(V):
                    self.devs[v][k] = dev_all[v]
        self._e2e_graph = None
        # what the graphed e2e step moves: the blocks (fields padded to 16 bytes) up, the views' losses down
        self.h2d_bytes = sum(b[0].numel() * b[0].element_size() for b in self.e2e_blocks)
        self.d2h_bytes = 4 * V

    def _init_flat(self, device):
        """Our training arm: the product's data-parallel plumbing (hugs_b200.dp).  ONE flat gradient buffer
        (FlatGradBuffer, gathered mode): the SH gradient (81 % of the bytes) is added by the backward kernel
        itself into its slice (ShGradSink), the small gradients are accumulated by autograd on the views'
        streams and copied into their slices right before the step's one allreduce; nothing is copied back."""
        from hugs_b200 import dp
        names = ["shs", "opacity", "s_means", "s_scales", "s_rots", "h_xyz", "h_rotmat", "h_scales", "h_weights"]
        self.buf = dp.FlatGradBuffer({n: getattr(self, n) for n in names if getattr(self, n) is not None},
                                     gather=True)
        self.sh_sink = self.buf.attach_sh_sink("shs")

    def _init_two_segments(self, wl, device, leaf):
        """Our training arm: the product's data-parallel plumbing (hugs_b200.dp).  Every parameter's .grad is
        a slice of ONE flat buffer (FlatGradBuffer); the parameters that feed the rasterizer directly have
        gradient sinks, so the backward kernels add into that buffer and the step's allreduce works on it as
        it is -- no torch.cat of inputs or gradients, no per-view AccumulateGrad kernels."""
        from hugs_b200 import dp
        e = lambda *shape: torch.zeros(*shape, device=device).requires_grad_(True)  # noqa: E731
        if self.n_human > 0:
            self.h_opacity, self.h_shs = leaf(wl["human"]["opacity"]), leaf(wl["human"]["shs"])
        else:
            self.h_opacity, self.h_shs = e(0, 1), e(0, 16, 3)
        if self.n_scene > 0:
            self.s_opacity, self.s_shs = leaf(wl["scene"]["opacity"]), leaf(wl["scene"]["shs"])
        else:
            self.s_means, self.s_scales, self.s_rots = e(0, 3), e(0, 3), e(0, 4)
            self.s_opacity, self.s_shs = e(0, 1), e(0, 16, 3)
        self.opacity = self.shs = None
        # sinks where the rows take 16-byte vector reductions (SH, quaternions) or one scalar (opacity); the
        # 12-byte rows of means3D / scales would be three scalar atomics each -- measured slower than a plain
        # store + AccumulateGrad (HB_SINK_ALL=1 switches them on for the experiment)
        sunk = ["h_shs", "s_shs", "h_opacity", "s_opacity", "s_rots"]
        if os.environ.get("HB_SINK_ALL", "1") == "1":
            sunk += ["s_means", "s_scales"]
        names = sunk + [n for n in ("s_means", "s_scales") if n not in sunk] + ["h_xyz", "h_rotmat", "h_scales", "h_weights"]
        params = {n: getattr(self, n) for n in names if getattr(self, n) is not None and getattr(self, n).numel() > 0}
        self.buf = dp.FlatGradBuffer(params)
        self.buf.attach_sinks([n for n in sunk if n in params])

    def leaves(self):
        if self.buf is not None:
            return dict(self.buf.params)
        names = ["h_xyz", "h_rotmat", "h_scales", "h_weights", "s_means", "s_scales", "s_rots", "opacity", "shs"]
        return {n: getattr(self, n) for n in names if getattr(self, n) is not None}

    def zero_grad(self):
        if not self.training or self.buf is not None:   # dp_step clears the flat buffer itself
            return
        for n, p in self.leaves().items():
            if n == "shs" and self.sh_sink is not None:
                if self.n_streams > 1:
                    self.sh_sink.zero_()  # every view adds (views run concurrently on several streams)
                else:
                    self.sh_sink.reset()  # the next backward pass overwrites the sink buffer
            else:
                p.grad = None
        self.means2D.grad = None

    def _gaussians(self, inp):
        """LBS of the human + the scene: the rasterizer's (means3D, scales, rotations)."""
        if self.n_human > 0:
            if "deformed" in inp:
                xyz, rotq, hsc = inp["deformed"]
            else:
                A = inp["A"].requires_grad_(True) if self.training else inp["A"]
                xyz, rotq, hsc = self.lbs_impl(A, self.h_xyz, self.h_weights, self.h_rotmat, self.h_scales, self.transl)
            if self.n_scene > 0:
                return torch.cat([xyz, self.s_means]), torch.cat([hsc, self.s_scales]), torch.cat([rotq, self.s_rots])
            return xyz, hsc, rotq
        return self.s_means, self.s_scales, self.s_rots

    def frame(self, inp):
        """One forward (+ backward when training) of one view.  Returns the loss tensor / the image (device)."""
        if not self.training:
            with torch.no_grad():
                return self._frame(inp)
        return self._frame(inp)

    def _frame(self, inp):
        out = self._forward(inp)
        if self.training:
            out.backward()
            return out.detach()
        return out

    def _forward(self, inp):
        """Forward of one view: the loss (with its graph) when training, else the image."""
        if self.two_seg:
            return self._forward_two_segments(inp)
        means3D, scales, rots = self._gaussians(inp)
        settings = self.module.GaussianRasterizationSettings(
            image_height=self.H, image_width=self.W, tanfovx=self.tanfovx, tanfovy=self.tanfovy, bg=inp["bg"],
            scale_modifier=1.0, viewmatrix=inp["viewmatrix"], projmatrix=inp["projmatrix"], sh_degree=3,
            campos=inp["campos"], prefiltered=False, debug=False)
        color_h = None
        if self.mode == "train_separate":
            color, color_h, radii = self.render_separate(
                settings, dict(means3D=means3D, means2D=self.means2D, opacities=self.opacity, shs=self.shs,
                               scales=scales, rotations=rots), self.n_human, self.human_bg)
        else:
            color, radii = self.module.GaussianRasterizer(settings)(
                means3D=means3D, means2D=self.means2D, opacities=self.opacity, shs=self.shs, scales=scales,
                rotations=rots)
        self.last_radii = radii
        self.last_node = color.grad_fn
        if not self.training:
            if "img_out" in inp:   # e2e: the image is the result, it goes back to the host
                inp["img_out"].copy_(color, non_blocking=True)
                inp["img_event"].record()
            return color
        if "G_ready" in inp:  # the upload of G runs on the copy stream; it is first needed here
            torch.cuda.current_stream(self.device).wait_event(inp["G_ready"])
        loss = (color * inp["G"]).sum()
        if color_h is not None:
            loss = loss + (color_h * inp["Gh"]).sum()
        if "loss_out" in inp:
            inp["loss_out"].copy_(loss.detach().reshape(1), non_blocking=True)
            inp["loss_event"].record()
        return loss

    def _forward_two_segments(self, inp):
        settings = self.module.GaussianRasterizationSettings(
            image_height=self.H, image_width=self.W, tanfovx=self.tanfovx, tanfovy=self.tanfovy, bg=inp["bg"],
            scale_modifier=1.0, viewmatrix=inp["viewmatrix"], projmatrix=inp["projmatrix"], sh_degree=3,
            campos=inp["campos"], prefiltered=False, debug=False)
        if self.n_human > 0:
            A = inp["A"].requires_grad_(True)
            xyz, rotq, hsc = self.lbs_impl(A, self.h_xyz, self.h_weights, self.h_rotmat, self.h_scales, self.transl)
        else:
            xyz, rotq, hsc = self.s_means[:0], self.s_rots[:0], self.s_scales[:0]
        # the screen-space dummy of gs_renderer.py:107-113: its values are never read, its .grad is per view
        means2D = torch.empty(self.P, 3, device=self.device, requires_grad=True)
        color, color_h, radii = self.render2(
            settings, (xyz, self.h_shs, self.h_opacity, hsc, rotq),
            (self.s_means, self.s_shs, self.s_opacity, self.s_scales, self.s_rots), means2D,
            self.human_bg if self.mode == "train_separate" else None)
        self.last_radii, self.last_node = radii, color.grad_fn
        if "G_ready" in inp:
            torch.cuda.current_stream(self.device).wait_event(inp["G_ready"])
        loss = (color * inp["G"]).sum()
        if color_h is not None:
            loss = loss + (color_h * inp["Gh"]).sum()
        if "loss_out" in inp:
            inp["loss_out"].copy_(loss.detach().reshape(1), non_blocking=True)
            inp["loss_event"].record()
        return loss

    def _settings(self, inp):
        return self.module.GaussianRasterizationSettings(
            image_height=self.H, image_width=self.W, tanfovx=self.tanfovx, tanfovy=self.tanfovy, bg=inp["bg"],
            scale_modifier=1.0, viewmatrix=inp["viewmatrix"], projmatrix=inp["projmatrix"], sh_degree=3,
            campos=inp["campos"], prefiltered=False, debug=False)

    def _dp_step(self, forward_of_view, after_view=None, exchange=True, prepass=False):
        """The product's data-parallel step (hugs_b200.dp.dp_step): clear the flat gradient buffer, this rank's
        views over the streams, backward of each, ONE allreduce of the flat buffer."""
        from hugs_b200 import dp
        self.means2D.grad = None

        def render_view(v):
            loss = forward_of_view(v)
            self.last_nodes[v], self.last_radii_v[v] = int(self.last_node.num_rendered), self.last_radii
            return loss
        return dp.dp_step(render_view, list(range(self.views)), self.buf, 0, 1,
                   streams=self.streams if self.n_streams > 1 else (), read_loss=False,
                          after_view=after_view, exchange=exchange,
                          pin_accumulators=os.environ.get("HB_PIN_ACC", "1") == "1",   # (0: tools/graph_bisect.py)
                          defer_scene=self.two_seg and os.environ.get("HB_DEFER", "1") == "1",
                          scene_prepass=self._scene_prepass if (prepass and self.two_seg and self.n_scene > 0
                                                                and os.environ.get("HB_PREPASS", "1") == "1") else None)

    def _scene_prepass(self):
        """The scene segment preprocessed for all of the step's (device-resident) cameras in one launch."""
        from hugs_b200 import segments
        return segments.scene_prepass([self._settings(dv) for dv in self.devs],
                                      (self.s_means, self.s_shs, self.s_opacity, self.s_scales, self.s_rots),
                                      self.n_human)

    def _deform_all(self, As):
        """animation: the rank's pose frames of the step through ONE batched LBS call when the
        implementation has one (ours), else frame by frame inside frame()."""
        if self.mode != "animation" or self.lbs_frames is None:
            return None
        with torch.no_grad():
            xyz, rotq, hsc = self.lbs_frames(As, self.h_xyz, self.h_weights, self.h_rotmat, self.h_scales,
                                             self.transl[None])
        return xyz, rotq, hsc

    def step_resident(self):
        """One step: every view of this rank's share of the view batch, gradients accumulated.  The two-segment
        path replays the local part of the step as one CUDA graph and does the gradient exchange after it."""
        if self.use_graph and self.buf is not None:
            self._replay()
            return
        self._step_resident_eager()

    def _step_resident_eager(self, exchange=True):
        if self.buf is not None:
            return self._dp_step(lambda v: self._forward(dict(self.devs[v])), exchange=exchange, prepass=True)
        self.zero_grad()
        if self.mode == "animation" and self.lbs_frames is not None:
            if not hasattr(self, "A_all"):
                self.A_all = torch.stack([dv["A"] for dv in self.devs])
            dall = self._deform_all(self.A_all)
        else:
            dall = None

        def inputs(v):
            inp = dict(self.devs[v])
            if dall is not None:
                inp["deformed"] = (dall[0][v], dall[1][v], dall[2][v])
            return inp

        if self.n_streams > 1:
            main = torch.cuda.current_stream(self.device)
            for s in self.streams:
                s.wait_stream(main)
            for v in range(self.views):
                with torch.cuda.stream(self.streams[v % self.n_streams]):
                    self.frame(inputs(v))
                self.last_nodes[v] = self.last_node
                self.last_radii_v[v] = self.last_radii
            for s in self.streams:
                main.wait_stream(s)
            return
        for v in range(self.views):
            self.frame(inputs(v))
            self.last_nodes[v] = self.last_node
            self.last_radii_v[v] = self.last_radii

    def _graphed_step(self):
        if self.graphed is None:
            from hugs_b200 import dp
            self.graphed = dp.GraphedStep(lambda: self._step_resident_eager(exchange=False))
            if self.buf.gather:   # the graph writes the tensors autograd adopted as .grad during the capture
                self._graph_grads = {k: p.grad for k, p in self.buf.params.items() if k not in self.buf.sunk}
        return self.graphed

    def _replay(self):
        out = self._graphed_step().replay()
        if self._graph_grads is not None:   # (an eager step in between -- the staged pass -- re-points .grad)
            for k, g in self._graph_grads.items():
                self.buf.params[k].grad = g
        self.buf.allreduce()
        return out

    def _prefetch(self):
        """The NEXT step's inputs, pinned host blocks -> device staging blocks, on the copy stream (behind the
        copy that emptied the staging blocks): runs while the current step is replayed."""
        g = self._e2e_graph
        with torch.cuda.stream(self.copy_stream):
            if g["free"] is not None:
                self.copy_stream.wait_event(g["free"])
            for host_b, stage_b, _ in self.e2e_blocks:
                stage_b.copy_(host_b, non_blocking=True)
            g["staged"].record(self.copy_stream)

    def _step_e2e_graphed(self):
        """End to end with the step replayed as a CUDA graph: the step's inputs arrive from pinned host memory in
        the staging blocks (uploaded on the copy stream during the previous step, the loader pattern: one
        upload per step, K uploads in K timed steps), are copied device-to-device to where the graph reads them
        (~0.06 ms for 200 MB), the graph is replayed, the gradients are exchanged, and the views' losses go to a
        pinned slot that the host reads two steps later."""
        if self._e2e_graph is None:
            self._e2e_graph = dict(free=None, staged=torch.cuda.Event(), k=0, done=[None, None],
                                   pinned=[torch.zeros(self.views).pin_memory() for _ in range(2)])
            self._prefetch()
        g = self._e2e_graph
        slot = g["k"] % 2
        if g["done"][slot] is not None:   # the losses of step k - 2
            g["done"][slot].synchronize()
            self.losses.extend(g["pinned"][slot].tolist())
        main = torch.cuda.current_stream(self.device)
        main.wait_event(g["staged"])
        with torch.no_grad():
            for _, stage_b, dev_b in self.e2e_blocks:
                dev_b.copy_(stage_b)
        g["free"] = torch.cuda.Event()
        g["free"].record(main)
        self._prefetch()
        losses = self._replay()
        g["pinned"][slot].copy_(torch.stack(losses), non_blocking=True)
        g["done"][slot] = torch.cuda.Event()
        g["done"][slot].record(main)
        g["k"] += 1

    def step_e2e(self):
        if self.use_graph and self.e2e_blocks is not None and self.buf is not None:
            return self._step_e2e_graphed()
        if self.buf is not None:
            def release(v):   # the view's backward has been enqueued: its G buffer may be refilled after it
                done = torch.cuda.Event()
                done.record()
                self.step_done[self._g_slot[v]] = done
            self._g_slot = {}
            self._small_dev = None
            if self.prefetch_small:
                if self._small_all is None:
                    recs = [torch.cat([t.reshape(-1) for k, t in h.items() if k not in ("G", "Gh", "_packed")])
                            for h in self.hosts]
                    self._small_len = int(recs[0].numel())
                    self._small_all = torch.stack(recs).pin_memory()
                self._small_dev = self._small_all.to(self.device, non_blocking=True)
            return self._dp_step(lambda v: self.frame_e2e(self.hosts[v], forward_only=True, view=v), release)
        self.zero_grad()
        dall = None
        if self.mode == "animation" and self.lbs_frames is not None:
            # the step's joint transforms travel as one pinned block
            if not hasattr(self, "A_host"):
                self.A_host = torch.stack([h["A"] for h in self.hosts]).pin_memory()
            dall = self._deform_all(self.A_host.to(self.device, non_blocking=True))
        if self.n_streams > 1:
            main = torch.cuda.current_stream(self.device)
            for s in self.streams:
                s.wait_stream(main)
            for v in range(self.views):
                with torch.cuda.stream(self.streams[v % self.n_streams]):
                    self.frame_e2e(self.hosts[v], None if dall is None else (dall[0][v], dall[1][v], dall[2][v]))
            for s in self.streams:
                main.wait_stream(s)
            return
        for v in range(self.views):
            self.frame_e2e(self.hosts[v], None if dall is None else (dall[0][v], dall[1][v], dall[2][v]))

    def frame_e2e(self, host, deformed=None, forward_only=False, view=None):
        """One frame with every per-frame input copied from pinned host memory (camera, joint
        transforms on the compute stream; G on the copy stream) and the result read back on the host:
        the loss when training, the IMAGE when rendering."""
        i = self.e2e_step % self.n_gbuf
        main = torch.cuda.current_stream(self.device)
        big = ("G", "Gh", "_packed")
        if self.training:
            with torch.cuda.stream(self.copy_stream):
                if self.step_done[i] is not None:  # buffer i was last read by the step before the previous one
                    self.copy_stream.wait_event(self.step_done[i])
                for k, buf in self.G_dev[i].items():
                    buf.copy_(host[k], non_blocking=True)
                self.G_ready[i].record(self.copy_stream)
        # camera + joint transforms travel as ONE pinned record (one async copy instead of five)
        small = [(k, v) for k, v in host.items() if k not in big and not (deformed is not None and k == "A")]
        if forward_only and getattr(self, "_small_dev", None) is not None:
            rec = self._small_dev[view]   # uploaded with the other views' records when the step began
        else:
            if "_packed" not in host:
                host["_packed"] = torch.cat([v.reshape(-1) for _, v in small]).pin_memory()
            rec = host["_packed"].to(self.device, non_blocking=True)
        inp, off = {}, 0
        for k, v in small:
            inp[k] = rec[off:off + v.numel()].view(v.shape)
            off += v.numel()
        if deformed is not None:
            inp["deformed"] = deformed
        # Read-back: every frame's result is copied to a pinned slot of a small ring and read by the host
        # when the slot comes round again (4 frames later; the host does not stall the device for the
        # 4-byte loss, and waits on PCIe for the 25 MB image when the copy is the bottleneck);
        # drain_e2e reads what is still pending before the closing event.
        s = self.e2e_step % 4
        ev = torch.cuda.Event()
        if self.training:
            if self.loss_ready[s] is not None:
                self.loss_ready[s].synchronize()
                self.losses.append(float(self.loss_host[s][0]))
            inp.update(self.G_dev[i])
            inp.update(G_ready=self.G_ready[i], loss_out=self.loss_host[s], loss_event=ev)
            out = self._forward(inp) if forward_only else self.frame(inp)
            self.loss_ready[s] = ev
        else:
            if self.img_ready[s] is not None:
                self.img_ready[s].synchronize()
                self.images_read += 1
            inp.update(img_out=self.img_host[s], img_event=ev)
            self.frame(inp)
            self.img_ready[s] = ev
        self.e2e_step += 1
        if self.training and forward_only:
            self._g_slot[view] = i   # released by _dp_step's after_view, once the backward is enqueued too
            return out
        done = torch.cuda.Event()
        done.record(main)
        self.step_done[i] = done
        return None

    def drain_e2e(self):
        """Read the results still pending (called inside the timed region, after the last step)."""
        g = getattr(self, "_e2e_graph", None)
        if g is not None:
            for j in (g["k"], g["k"] + 1):   # step k - 2, then step k - 1
                if g["done"][j % 2] is not None:
                    g["done"][j % 2].synchronize()
                    g["done"][j % 2] = None
                    self.losses.extend(g["pinned"][j % 2].tolist())
        for s in range(4):
            if self.loss_ready[s] is not None:
                self.loss_ready[s].synchronize()
                self.loss_ready[s] = None
                self.losses.append(float(self.loss_host[s][0]))
            if self.img_ready[s] is not None:
                self.img_ready[s].synchronize()
                self.img_ready[s] = None
                self.images_read += 1

    def allreduce_grads(self):
        """The gradient exchange is part of the product's step (hugs_b200.dp.dp_step -> FlatGradBuffer.allreduce,
        inside _dp_step); the reference arm is single-GPU (the reference has no multi-GPU path)."""
        return


def our_impl():
    import diff_gaussian_rasterization as dgr
    import hugs_b200
    from hugs_b200 import layer, segments

    def lbs_impl(A, xyz, w, rotmat, scales, transl):
        return hugs_b200.deform_gaussians(A, xyz, w, rotmat, scales, smpl_scale=1.0, transl=transl)

    def lbs_frames(As, xyz, w, rotmat, scales, transl):
        return hugs_b200.lbs_deform_frames(As, xyz, w, rotmat, scales, transl.expand(As.shape[0], 3).contiguous(), 1.0)

    def separate(settings, t, n_human, human_bg):
        # ONE rasterizer pass: the human image comes from the joint pass's sorted lists (hugs_b200/layer.py)
        return layer.rasterize_with_layer(t["means3D"], t["means2D"], t["opacities"], t["scales"], t["rotations"],
                                          settings, n_human, human_bg, shs=t["shs"])

    def render2(settings, seg0, seg1, means2D, human_bg):
        # two input segments (no torch.cat); with `human_bg` also the separate human image, same pass
        if human_bg is not None:
            return layer.rasterize_two_segments_with_layer(seg0, seg1, means2D, settings, human_bg)
        color, radii = segments.rasterize_two_segments(seg0, seg1, means2D, settings)
        return color, None, radii

    return dict(module=dgr, lbs=lbs_impl, lbs_frames=lbs_frames, separate=separate, render2=render2)


def reference_impl():
    """The unmodified reference: CUDA rasterizer build + PyTorch LBS chain (oracle/_ref)."""
    import refutil

    ref = refutil.load_ref_rasterizer()
    lbs, rot = refutil.load_ref_lbs()

    def lbs_impl(A, xyz, w, rotmat, scales, transl):
        pose = torch.zeros(1, 72, device=xyz.device)
        # hugs/models/hugs_trimlp.py:470-518 (smpl_scale = 1)
        verts, _, T, _, _ = lbs.lbs_extra(A[None], xyz[None], None, w, pose, disable_posedirs=True, pose2rot=True)
        dxyz = verts.squeeze(0) + transl[None]
        T = T.squeeze(0)
        return dxyz, rot.matrix_to_quaternion(T[:, :3, :3] @ rotmat), scales * 1.0

    def separate(settings, t, n_human, human_bg):
        # hugs/renderer/gs_renderer.py:56-82: render(everything), then render(the human alone)
        color, radii = ref.GaussianRasterizer(settings)(**t)
        h = {k: v[:n_human] for k, v in t.items() if k != "means2D"}
        h["means2D"] = torch.zeros(n_human, 3, device=color.device, requires_grad=True)
        color_h, _ = ref.GaussianRasterizer(settings._replace(bg=human_bg))(**h)
        return color, color_h, radii

    return dict(module=ref, lbs=lbs_impl, lbs_frames=None, separate=separate)


# ------------------------------------------------------------------------------------------------
# roofline accounting (algorithmic bytes: SURVEY.md 8(d) / DESIGN.md "Algorithmic bytes")
# ------------------------------------------------------------------------------------------------
def stage_bytes(P, visible, R, W, H, n_human, sort_bits, views=1, sh_sink=False, batched_scene=False, R1=None):
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    culled = P - visible
    passes = (sort_bits + 7) // 8
    out = _stage_bytes(P, visible, culled, R, W, H, tiles, passes, n_human, views, sh_sink)
    if R1 is not None:
        # The two-level multisplit (csrc/binning2.cu) never sorts the R pairs: `sort_pairs` is the argsort of the P
        # depth words (histogram + four 8-bit passes over key + index) plus ONE 8-bit pass over the R1 level-1
        # (Gaussian, super-tile) entries; `duplicate_with_keys` emits those entries.
        out["sort_pairs"] = 8 * P + 4 * 16 * P + 24 * R1
        out["duplicate_with_keys"] = 24 * P + 8 * R1
    if batched_scene and P > n_human:
        # The scene segment's per-Gaussian passes run ONCE per step for all `views` views (csrc/preprocess.cu:
        # preprocess_views_kernel, csrc/preprocess_backward.cu: scene_backward_views_kernel): per frame, the
        # human rows as before plus the scene's once-per-step bytes divided by the views and its per-view bytes.
        Ps = P - n_human
        vis_s = visible * Ps // P
        vis_h, cul_s = visible - vis_s, Ps - vis_s
        cul_h = n_human - vis_h
        out["preprocess"] = (311 * vis_h + 20 * cul_h) + 236 * Ps // views + 75 * vis_s + 20 * cul_s
        out["preprocess_backward"] = ((751 * vis_h + 68 * cul_h) + (236 + 2 * 236) * Ps // views + 41 * vis_s + 4 * Ps
                                      + 24 * Ps)   # (the last term: the per-view unpack of the screen-space gradient)
    return out


def _stage_bytes(P, visible, culled, R, W, H, tiles, passes, n_human, views, sh_sink):
    return {
        "lbs_forward": 172 * n_human,
        "preprocess": 311 * visible + 20 * culled,
        "inclusive_scan": 8 * P,
        "duplicate_with_keys": 24 * P + 12 * R,
        "sort_pairs": (8 + 24 * passes) * R,
        "identify_tile_ranges": 8 * R + 8 * tiles,
        # level 2 of the two-level binning (csrc/binning2.cu): the compulsory output -- the sorted list and
        # the ranges; it replaces the per-pair sort passes and identifyTileRanges of the reference structure
        "tile_split": 4 * R + 8 * tiles,
        "render_forward": 40 * R + 20 * W * H + 8 * tiles,
        "render_backward": 76 * R + 20 * W * H,
        # per view.  With the SH gradient sink, views 2..V of a step ADD their 192-B SH gradient row to the
        # sink (read + write, visible Gaussians only) instead of writing it: 559 + 192 B per visible
        # Gaussian, 68 B (the zero rows of the other outputs) per culled one.
        "preprocess_backward": (559 * visible + 260 * culled if not sh_sink else
                                ((559 * visible + 260 * culled) + (views - 1) * (751 * visible + 68 * culled)) // views),
        "lbs_backward": 352 * n_human,
    }


def timed_loop(fn, steps, device, barrier, finish=None):
    """EXACTLY `steps` calls of fn between barrier+synchronize brackets, CUDA-event timed.
    The cyclic garbage collector is paused for the timed region (both arms): a generation-2 sweep of
    the interpreter's ~10^6 live objects takes 20-30 ms, ten frames' worth, whenever it happens to fire."""
    import gc
    gc.collect()
    gc.disable()
    try:
        barrier()
        torch.cuda.synchronize(device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        if finish is not None:
            finish()
        e1.record()
        torch.cuda.synchronize(device)
        barrier()
    finally:
        gc.enable()
    return e0.elapsed_time(e1)  # ms


def cpu_oracle_frames(wl, seconds, min_frames=2):
    """Time the CPU oracle port (oracle/splat_oracle.c, OpenMP) on full frames of the workload for
    about `seconds`.  Returns (frames/s, threads, frames run)."""
    from oracle import oracle as O
    import numpy as np

    cam = wl["cam"]
    n = lambda t: t.detach().cpu().numpy()  # noqa: E731
    W, H = wl["W"], wl["H"]
    threads = os.cpu_count() or 1
    training = wl["mode"] in ("train", "train_separate")

    def one():
        parts = []
        if wl["n_human"] > 0:
            h = wl["human"]
            xyz, q, hs, _ = O.lbs_forward(n(wl["A"]), n(h["xyz"]), n(h["lbs_weights"]), n(h["rotmat"]), n(h["scales"]),
                                          n(h["transl"][0]), 1.0)
            parts.append((xyz, hs, q, n(h["opacity"]), n(h["shs"])))
        if wl["n_scene"] > 0:
            s = wl["scene"]
            parts.append((n(s["means3D"]), n(s["scales"]), n(s["rotations"]), n(s["opacity"]), n(s["shs"])))
        means, scales, rots, opac, shs = (np.concatenate([p[i] for p in parts]) for i in range(5))
        kw = dict(shs=shs, degree=3, scales=scales, rotations=rots)
        fw = O.forward(means, opac, n(cam["viewmatrix"]), n(cam["projmatrix"]), n(cam["campos"]), n(cam["bg"]), W, H,
                       cam["tanfovx"], cam["tanfovy"], **kw)
        if training:
            O.backward(fw, n(wl["G"]), means, n(cam["viewmatrix"]), n(cam["projmatrix"]), n(cam["campos"]),
                       n(cam["bg"]), W, H, cam["tanfovx"], cam["tanfovy"], **kw)

    one()  # warm-up (page-in, OpenMP pool)
    t0 = time.perf_counter()
    frames = 0
    while frames < min_frames or (time.perf_counter() - t0) < seconds:
        one()
        frames += 1
        if frames >= 200:
            break
    dt = time.perf_counter() - t0
    what = "LBS fwd + raster fwd + bwd" if training else "LBS fwd + raster fwd"
    return frames / dt, threads, "%d full frames of %s (%s) on the C/OpenMP oracle" % (frames, wl["name"], what)


def cpu_torch_frames(wl, seconds, min_frames=1):
    """BASELINE.json configs[0]: the pure-PyTorch CPU evaluation of the reference math (oracle/torch_cpu.py:
    the reference's own lbs_extra / matrix_to_quaternion / eval_sh + torch restatements of the CUDA-only
    preprocess and tile composite), forward + autograd backward, on the host cores."""
    from oracle import torch_cpu as TC
    h, cam = wl["human"], wl["cam"]
    names = ("xyz", "rotmat", "scales", "lbs_weights", "opacity", "shs")
    threads = torch.get_num_threads()

    def one():
        t = {k: h[k].detach().clone().requires_grad_(True) for k in names}
        A = wl["A"].clone().requires_grad_(True)
        xyz, q, sc = TC.deform(A, t["xyz"], t["lbs_weights"], t["rotmat"], t["scales"], h["transl"][0])
        img = TC.render(xyz, sc, q, t["opacity"], t["shs"], 3, cam["viewmatrix"], cam["projmatrix"], cam["campos"],
                        cam["bg"], wl["W"], wl["H"], cam["tanfovx"], cam["tanfovy"])
        (img * wl["G"]).sum().backward()

    one()
    t0 = time.perf_counter()
    frames = 0
    while frames < min_frames or (time.perf_counter() - t0) < seconds:
        one()
        frames += 1
    dt = time.perf_counter() - t0
    return frames / dt, threads, ("%d full frames of %s (reference lbs_extra + eval_sh, torch preprocess + tile "
                                  "composite, autograd backward) in pure PyTorch on the CPU" % (frames, wl["name"]))


def cpu_baseline_of(name, seconds, min_frames=2):
    wl = make_workload(name, 0, 1, "cpu")
    wl["name"] = name
    if WORKLOADS[name].get("cpu") == "torch":
        fps, threads, sample = cpu_torch_frames(wl, seconds)
    else:
        fps, threads, sample = cpu_oracle_frames(wl, seconds, min_frames)
    return {"value": round(fps, 4), "unit": UNIT, "cores": threads, "kind": "port", "sample": sample}


# ------------------------------------------------------------------------------------------------
# main
# ------------------------------------------------------------------------------------------------
def metric_of(name):
    w = WORKLOADS[name]
    what = {"train": "fwd+bwd", "train_separate": "fwd+bwd (joint + separate human image)", "forward": "forward-only",
            "animation": "LBS + forward-only"}[w["mode"]]
    if name == "human_scene_1080p":
        return METRIC
    return "%s frames/s, %s" % (what, name)


def main():
    args = parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != max(args.gpus, 1) and world > 1:
        print("warning: --gpus %d but WORLD_SIZE=%d" % (args.gpus, world), file=sys.stderr)
    n_gpus = world
    w = WORKLOADS[args.workload]
    n_human, n_scene, W, H, mode = w["n_human"], w["n_scene"], w["W"], w["H"], w["mode"]
    training = mode in ("train", "train_separate")
    scaling = args.scaling
    if mode == "animation":
        scaling = "strong"       # the step is the workload's `frames` pose frames, whatever N is
    elif scaling == "strong":
        gb = args.global_batch or args.views_per_rank
        if gb % n_gpus:
            raise SystemExit("--global-batch %d is not a multiple of %d ranks" % (gb, n_gpus))
        args.views_per_rank = gb // n_gpus
    frames_per_step = w["frames"] if mode == "animation" else args.views_per_rank * n_gpus
    config = {"workload": args.workload, "n_human": n_human, "n_scene": n_scene, "image": "%dx%d" % (W, H),
              "sh_degree": 3, "mode": mode,
              "views_per_rank_per_step": (frames_per_step / n_gpus if mode == "animation" else args.views_per_rank),
              "global_view_batch": frames_per_step,
              "parallelism": ("dp%d (view batch split round-robin across ranks, gradients accumulated over a rank's "
                              "views, one allreduce of the parameter gradients per step)" % n_gpus) if training else
                             ("frames split round-robin over %d rank(s), no collective" % n_gpus),
              "l2": "inputs larger than L2 (per-frame working set > 1 GB vs 126 MB L2), no flush"}
    if H * W < 1920 * 1080:
        config["l2"] = ("per-frame working set ~%d MB, below the 126 MB L2: successive views use different cameras, "
                        "joint transforms and loss gradients, no flush" % ((n_human + n_scene) * 300 // 1_000_000))

    # ---- reference arm -------------------------------------------------------------------------
    if args.impl == "reference":
        if rank != 0:
            return 0
        return run_reference(args, config)

    if not torch.cuda.is_available():
        raise SystemExit("bench.py (our arm) needs a CUDA device: the product has no CPU path")
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if world > 1:
            dist.barrier()

    from hugs_b200 import _native
    V = args.views_per_rank
    wl = make_workload(args.workload, rank, world, device, views=V)
    rep = Replica(wl, device, our_impl(), sh_sink=True)
    V = rep.views        # animation: this rank's share of the step's frames

    def step_resident():
        rep.step_resident()
        rep.allreduce_grads()

    def step_e2e():
        rep.step_e2e()
        rep.allreduce_grads()

    use_graph, rep.use_graph = rep.use_graph, False
    step_resident()   # one eager step: the pair counts below come from it (a captured step reports capacities)
    torch.cuda.synchronize(device)
    # pair / visible counts: mean over this rank's views (they size the per-stage algorithmic bytes)
    R1 = _native.last_pair_counts()[1] or None   # level-1 entries of the last view's forward call (default binning mode)
    visible = int(round(sum(int((r > 0).sum()) for r in rep.last_radii_v) / V))
    if training:
        R = int(round(sum(int(nd if isinstance(nd, int) else nd.num_rendered) for nd in rep.last_nodes) / V))
    else:   # rendering runs under no_grad (no autograd node): count the pairs of this rank's first view once
        from hugs_b200 import rasterizer as _rz
        inp = dict(rep.devs[0])
        with torch.no_grad():
            m3, sc_, rq = rep._gaussians(inp)
        R = _rz.rasterize_gaussians_forward(
            inp["bg"], m3, torch.Tensor([]), rep.opacity, sc_, rq, 1.0, torch.Tensor([]), inp["viewmatrix"],
            inp["projmatrix"], rep.tanfovx, rep.tanfovy, H, W, rep.shs, 3, inp["campos"], False, False)[0]
        del m3, sc_, rq

    _native.launch_count(1)
    step_resident()
    torch.cuda.synchronize(device)
    launches_per_step = int(_native.launch_count(1))   # counted on an eager step: a replay launches the same kernels
    rep.use_graph = use_graph
    for _ in range(max(args.warmup, 3)):
        step_resident()
    torch.cuda.synchronize(device)

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    ms = timed_loop(step_resident, args.steps, device, barrier)
    launches = launches_per_step * args.steps
    for _ in range(2):
        step_e2e()
    rep.drain_e2e()
    rep.losses.clear()
    rep.images_read = 0
    ms_e2e = timed_loop(step_e2e, args.steps, device, barrier, finish=rep.drain_e2e)
    if training:
        assert len(rep.losses) == args.steps * V and all(math.isfinite(x) for x in rep.losses), \
            "e2e: a loss was not read back"
    else:
        assert rep.images_read == args.steps * V, "e2e: an image was not read back"
    clocks = sampler.stop() if sampler else None

    # per-stage device times: same K steps again with the library's stage events switched on
    _native.read_stage_times(reset=True)
    # (one stream in this pass: a stage's events must not span the other stream's kernels)
    _native.stage_timing_enable(1)
    rep.comm_events = []
    if rep.buf is not None:
        rep.buf.exchange_events = rep.comm_events
    n_streams, rep.n_streams = rep.n_streams, 1
    use_graph, rep.use_graph = rep.use_graph, False   # the stage events are not part of a captured step
    ms_staged = timed_loop(step_resident, args.steps, device, barrier)
    rep.n_streams, rep.use_graph = n_streams, use_graph
    _native.stage_timing_enable(0)
    stage_ms = _native.read_stage_times(reset=True)
    # data-parallel accounting (this pass only): time inside the gradient exchange per step -- waiting for
    # the slowest rank included -- on every rank
    comm_ms = sum(a.elapsed_time(b) for a, b in rep.comm_events) / max(len(rep.comm_events), 1)
    rep.comm_events = None
    if rep.buf is not None:
        rep.buf.exchange_events = None
    dp_info = None
    if world > 1 and training:
        t = torch.tensor([comm_ms, ms_staged / args.steps], dtype=torch.float64, device=device)
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        dp_info = {"exchange_ms_per_step_by_rank": [round(float(x[0]), 3) for x in allt],
                   "exchange_ms_min": round(min(float(x[0]) for x in allt), 3),
                   "gradient_bytes": int(sum(p.numel() * 4 for p in rep.leaves().values())),
                   "collectives_per_step": 1,
                   "through": "hugs_b200.dp.dp_step / FlatGradBuffer.allreduce (the product's API)",
                   "note": "min over ranks ~ pure transfer (the last rank to arrive does not wait); the rest is "
                           "waiting for the slowest rank's views"}

    from hugs_b200 import dp
    ms = dp.max_over_ranks(ms, device)
    ms_e2e = dp.max_over_ranks(ms_e2e, device)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peak, peak_kind = peaks()
    sort_bits = int(_native.sort_bits(W, H))
    nbytes = stage_bytes(rep.P, visible, R, W, H, n_human, sort_bits, views=V, sh_sink=(rep.sh_sink is not None or rep.buf is not None),
                         batched_scene=rep.two_seg and os.environ.get("HB_DEFER", "1") == "1"
                         and os.environ.get("HB_PREPASS", "1") == "1", R1=R1)
    stages = {}
    for name, (tot, calls) in stage_ms.items():
        if calls == 0:
            continue
        per = tot / (args.steps * V)  # per frame; a stage may be entered more than once (the two sort levels)
        gbs = nbytes[name] / (per * 1e-3) / 1e9
        stages[name] = {"ms": round(per, 4), "bytes": int(nbytes[name]), "gbs": round(gbs, 1),
                        "frac": round(gbs / peak, 4)}
    dominant = max(stages, key=lambda k: stages[k]["ms"]) if stages else None
    roofline = None
    # DRAM bytes per launch + issue-slot utilisation from the committed ncu --set full capture of the
    # headline workload (null for the others)
    traffic, traffic_src, issue_pct, issue_src = None, None, {}, None
    try:
        if args.workload in ("human_scene_1080p", "human_scene_separate_1080p"):
            with open(os.path.join(ROOT, "profiles", "ncu_dram_traffic.json")) as f:
                tj = json.load(f)
            traffic, traffic_src = tj["bytes_per_frame"], tj["source"]
            issue_pct, issue_src = tj.get("issue_active_pct", {}), tj.get("issue_source")
            for name in stages:
                if name in traffic and args.workload == "human_scene_1080p":
                    stages[name]["dram_traffic"] = int(traffic[name])
                if name in issue_pct:
                    stages[name]["issue_active_pct_ncu"] = issue_pct[name]
    except Exception:
        traffic = None
    if dominant:
        s = stages[dominant]
        issue_bound = dominant in ("render_forward", "render_backward")
        roofline = {"kernel": dominant, "bound": "issue" if issue_bound else "hbm",
                    "achieved": s["gbs"], "peak": peak, "unit": "GB/s", "frac": s["frac"],
                    "traffic": int(traffic[dominant]) if traffic and dominant in traffic else None,
                    "traffic_source": traffic_src,
                    "issue_frac": round(issue_pct[dominant] / 100.0, 4) if dominant in issue_pct else None,
                    "issue_source": issue_src,
                    "limiter": ("instruction issue: the blend kernels keep the warp schedulers' issue slots busy "
                                "(`issue_frac`, ncu) while their DRAM traffic (`traffic`) is a small fraction of the "
                                "algorithmic gather bytes -- a tile's Gaussians are served by L2; `achieved`/`frac` "
                                "are the HBM figures on the algorithmic bytes of SURVEY.md 8(d), kept for comparison "
                                "(DESIGN.md section 5); the HBM-bound stages are in `stages`") if issue_bound else "hbm",
                    "peak_source": peak_kind,
                    "ms_per_launch": s["ms"], "algorithmic_bytes": s["bytes"],
                    "timed": "per-stage CUDA events on the launching stream, %d frames on ONE stream (%.3f ms/frame "
                             "with events on; the timed region above alternates views over %d streams)"
                             % (args.steps * V, ms_staged / (args.steps * V), rep.n_streams)}
    frame_bytes = sum(nbytes[k] for k in stages)
    survey_bytes = (902 * rep.P + 288 * R + 40 * W * H) if training else frame_bytes
    total_frames = frames_per_step * args.steps
    line = {
        "metric": metric_of(args.workload), "value": round(total_frames / (ms * 1e-3), 3), "unit": UNIT,
        "n_gpus": n_gpus, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms / args.steps, 4),
        "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(config, P=rep.P, R=R, visible=visible, sort_bits=sort_bits),
        "impl": {"name": "ours", "cuda_streams_per_rank": rep.n_streams,
                 "batched_scene_passes": bool(rep.two_seg and os.environ.get("HB_DEFER", "1") == "1"),
                 "cuda_graph": ("value: the rank's local step is one CUDA-graph replay (%s), allreduce outside it; "
                                "e2e: inputs prefetched one step ahead on a copy stream into staging blocks, "
                                "device-to-device copy, the same graph replay, losses read two steps later"
                                % ("views over %d streams" % rep.n_streams))
                 if rep.use_graph else None},
        "e2e": {"value": round(total_frames / (ms_e2e * 1e-3), 3), "unit": UNIT,
                "h2d_bytes_per_step": int(rep.h2d_bytes), "d2h_bytes_per_step": int(rep.d2h_bytes),
                "result_read_back": "loss (4 B) per view" if training else "the rendered image (3xHxW f32) per frame",
                "ms_per_step": round(ms_e2e / args.steps, 4)},
        "gpu_launches": launches,
        "dp": dp_info,
        "clocks": clocks,
        "roofline": roofline,
        "stages": stages,
        # `algorithmic_bytes`: SURVEY.md 8(d)'s per-frame model of the REFERENCE's structure (902 P + 288 R + 40 W H),
        # the yardstick every round is measured on; `own_structure_bytes`: the sum of `stages[*].bytes`, what this
        # implementation's structure needs (no sort over the pairs, scene passes once per step)
        "frame": {"algorithmic_bytes": int(survey_bytes),
                  "ms": round(ms * n_gpus / total_frames, 4),
                  "hbm_frac": round(survey_bytes / (ms * n_gpus / total_frames * 1e-3) / 1e9 / peak, 4),
                  "own_structure_bytes": int(frame_bytes),
                  "own_structure_hbm_frac": round(frame_bytes / (ms * n_gpus / total_frames * 1e-3) / 1e9 / peak, 4)},
    }
    if n_gpus == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_of(args.workload, args.cpu_baseline_seconds)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_reference(args, config):
    """Reference arm: the unmodified reference through its own public API (rank 0 only)."""
    import refutil

    w = WORKLOADS[args.workload]
    training = w["mode"] in ("train", "train_separate")
    base = {"impl": "reference", "metric": metric_of(args.workload), "unit": UNIT, "n_gpus": 1, "steps": args.steps,
            "warmup": max(args.warmup, 3), "higher_is_better": True, "scaling": config.get("scaling", "weak"),
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config}
    base["scaling"] = "strong" if w["mode"] == "animation" or args.scaling == "strong" else "weak"
    if torch.cuda.is_available() and refutil.ref_available():
        device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
        torch.cuda.set_device(device)
        V = args.views_per_rank
        wl = make_workload(args.workload, 0, 1, device, views=V)
        rep = Replica(wl, device, reference_impl())
        V = rep.views
        for _ in range(max(args.warmup, 3)):
            rep.step_resident()
        sampler = ClockSampler(device.index or 0)
        sampler.start()
        ms = timed_loop(rep.step_resident, args.steps, device, lambda: None)
        for _ in range(2):
            rep.step_e2e()
        rep.drain_e2e()
        rep.losses.clear()
        rep.images_read = 0
        ms_e2e = timed_loop(rep.step_e2e, args.steps, device, lambda: None, finish=rep.drain_e2e)
        assert (len(rep.losses) if training else rep.images_read) == args.steps * V, "e2e: a result was not read back"
        clocks = sampler.stop()
        v = V * args.steps / (ms * 1e-3)
        ve = V * args.steps / (ms_e2e * 1e-3)
        base.update({
            "value": round(v, 3), "ms_per_step": round(ms / args.steps, 4), "clocks": clocks,
            "e2e": {"value": round(ve, 3), "unit": UNIT, "h2d_bytes_per_step": int(rep.h2d_bytes),
                    "d2h_bytes_per_step": int(rep.d2h_bytes), "ms_per_step": round(ms_e2e / args.steps, 4)},
            "cpu_baseline": {"value": round(ve, 3), "unit": UNIT, "cores": 0, "kind": "reference",
                             "sample": "the reference has no CPU rasterizer: this is its own CUDA build "
                                       "(oracle/_ref, unmodified sources, sm_100) + its PyTorch LBS chain on the "
                                       "same B200, full frames of " + args.workload},
            "reference_device": "cuda",
        })
        base["config"] = dict(config, P=rep.P)
        if training:
            base["config"]["R"] = int(round(sum(int(nd.num_rendered) for nd in rep.last_nodes) / V))
        print(json.dumps(base), flush=True)
        return 0
    # no GPU or no reference build: the CPU oracle port on the host cores
    cb = cpu_baseline_of(args.workload, max(10.0, args.cpu_baseline_seconds), min_frames=max(args.steps, 2))
    fps = cb["value"]
    base.update({
        "value": fps, "ms_per_step": round(1e3 / fps, 3),
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "cpu_baseline": cb,
        "reference_device": "cpu",
    })
    print(json.dumps(base), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
