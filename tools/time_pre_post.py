"""GPU-box timing of the two widening rows either side of the path:
  * tri-plane sampling, forward and forward+backward, 32 features, 256^3, 110 210 points, against the
    reference's three F.grid_sample calls + cat on the same GPU.  Algorithmic HBM bytes: forward
    N*(12 + 3C*4) + the planes once; backward N*(12 + 3C*4 + 12) + plane gradients zeroed and written
    (2x) + the planes once.
  * densification statistics on 1 M Gaussians against the reference's boolean-mask lines.
Prints one JSON line per op."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "ml-hugs_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
import torch.nn.functional as F

from hugs_b200 import renderer
from hugs_b200.triplane import TriPlane


def timeit(fn, iters=50, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6551.4


def triplane():
    n, C, R = 110210, 32, 256
    tp = TriPlane(check_range=False).cuda()
    x = (torch.rand(n, 3, device="cuda") * 2 - 1).requires_grad_(True)
    w = torch.randn(n, 3 * C, device="cuda")
    planes = [p.detach().clone().contiguous().requires_grad_(True) for p in tp.parameters()]

    def ours_f():
        with torch.no_grad():
            return tp(x)

    def ours_fb():
        x.grad = None
        for p in tp.parameters():
            p.grad = None
        tp(x).backward(w)

    def ref_fwd():
        c = (((x - 0.0) / 2.0 + 0.5) * 2 - 1).reshape(1, -1, 1, 3)
        return torch.cat([F.grid_sample(pl, c[..., idx], align_corners=True)[0, :, :, 0].transpose(0, 1)
                          for pl, idx in zip(planes, ([0, 1], [0, 2], [1, 2]))], dim=1)

    def ref_f():
        with torch.no_grad():
            return ref_fwd()

    def ref_fb():
        x.grad = None
        for p in planes:
            p.grad = None
        ref_fwd().backward(w)

    # kernels alone, straight through the C ABI (no autograd, no allocation)
    import ctypes
    from hugs_b200 import _native as N
    tex = [p.detach()[0].permute(1, 2, 0) for p in tp.parameters()]
    assert all(v.is_contiguous() for v in tex)
    xd, out = x.detach(), torch.empty(n, 3 * C, device="cuda")
    gtex = [torch.zeros_like(v) for v in tex]
    gx = torch.empty(n, 3, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def k_f():
        N.triplane_forward(n, 32, R, R, R, tex[0].data_ptr(), tex[1].data_ptr(), tex[2].data_ptr(), xd.data_ptr(),
                           0.0, 2.0, 1e-3, out.data_ptr(), None, st)

    def k_b():
        N.triplane_backward(n, 32, R, R, R, tex[0].data_ptr(), tex[1].data_ptr(), tex[2].data_ptr(), xd.data_ptr(),
                            0.0, 2.0, w.data_ptr(), gtex[0].data_ptr(), gtex[1].data_ptr(), gtex[2].data_ptr(),
                            gx.data_ptr(), st)

    t = {k: timeit(f) for k, f in (("ours_f", ours_f), ("ours_fb", ours_fb), ("ref_f", ref_f), ("ref_fb", ref_fb),
                                   ("k_f", k_f), ("k_b", k_b))}
    plane_bytes = 3 * C * R * R * 4
    fwd_bytes = n * (12 + 3 * C * 4) + plane_bytes
    peak = peak_gbs()
    kb_bytes = n * (12 + 3 * C * 4 + 12) + 2 * plane_bytes      # kernel alone: no zero fill
    print(json.dumps({"op": "triplane 32x256^3, 110210 points", "ours_fwd_ms": round(t["ours_f"], 4),
                      "ours_fwd_bwd_ms": round(t["ours_fb"], 4), "torch_reference_fwd_ms": round(t["ref_f"], 4),
                      "torch_reference_fwd_bwd_ms": round(t["ref_fb"], 4),
                      "speedup_fwd": round(t["ref_f"] / t["ours_f"], 2),
                      "speedup_fwd_bwd": round(t["ref_fb"] / t["ours_fb"], 2),
                      "fwd_kernel_ms": round(t["k_f"], 4), "bwd_kernel_ms": round(t["k_b"], 4),
                      "fwd_algorithmic_bytes": fwd_bytes, "fwd_hbm_frac": round(fwd_bytes / t["k_f"] / 1e6 / peak, 4),
                      "bwd_algorithmic_bytes": kb_bytes, "bwd_hbm_frac": round(kb_bytes / t["k_b"] / 1e6 / peak, 4)}))


def stats():
    n = 1_000_000
    radii = torch.randint(-20, 40, (n,), device="cuda", dtype=torch.int32).clamp_min(0)
    grad = torch.randn(n, 3, device="cuda") * 1e-3
    acc, den, mx = torch.zeros(n, 1, device="cuda"), torch.zeros(n, 1, device="cuda"), torch.zeros(n, device="cuda")

    def ours():
        renderer.add_densification_stats(radii, grad, acc, den, mx)

    def ref():
        vis = radii > 0
        mx[vis] = torch.max(mx[vis], radii[vis])
        acc[vis] += torch.norm(grad[:vis.shape[0]][vis, :2], dim=-1, keepdim=True)
        den[vis] += 1

    t_o, t_r = timeit(ours), timeit(ref)
    vis = int((radii > 0).sum())
    nbytes = 4 * n + vis * (8 + 24)
    print(json.dumps({"op": "densification stats, 1M Gaussians, %d visible" % vis, "ours_ms": round(t_o, 4),
                      "torch_reference_ms": round(t_r, 4), "speedup": round(t_r / t_o, 2),
                      "algorithmic_bytes": nbytes, "hbm_frac": round(nbytes / t_o / 1e6 / peak_gbs(), 4)}))


def two_segments():
    """render_human_scene(human_scene) forward + backward at the headline size: the two-segment entry
    points against the reference's structure (five torch.cat + the rasterizer on the concatenation)."""
    import scenes
    n_h, n_s = 110210, 1000000
    sc = scenes.make_scene(n_s, n_h, 1920, 1080, 1400.0, seed=1002, device="cuda")
    data = dict(image_height=1080, image_width=1920, fovx=sc["fovx"], fovy=sc["fovy"],
                world_view_transform=sc["viewmatrix"], full_proj_transform=sc["projmatrix"],
                camera_center=sc["campos"])
    names = dict(xyz="means3D", shs="shs", opacity="opacity", scales="scales", rotq="rotations")

    def gs(lo, hi):
        o = {k: sc[v][lo:hi].detach().clone().requires_grad_(True) for k, v in names.items()}
        o["active_sh_degree"] = 3
        return o
    human, scene = gs(0, n_h), gs(n_h, n_h + n_s)
    G = sc["dL_dcolor"]

    def clear():
        for o in (human, scene):
            for k in names:
                o[k].grad = None

    def seg():
        clear()
        pkg = renderer.render_human_scene(data, human, scene, sc["bg"])
        (pkg["render"] * G).sum().backward()

    def cat():
        clear()
        args = {a: torch.cat([human[k], scene[k]], dim=0) for k, a in renderer._KEYS}
        pkg = renderer.render(data=data, bg_color=sc["bg"], active_sh_degree=3, **args)
        (pkg["render"] * G).sum().backward()

    t_seg, t_cat = timeit(seg, iters=20), timeit(cat, iters=20)
    print(json.dumps({"op": "render_human_scene fwd+bwd, 110210 human + 1000000 scene @1080p",
                      "two_segments_ms": round(t_seg, 4), "torch_cat_ms": round(t_cat, 4),
                      "saved_ms": round(t_cat - t_seg, 4), "copy_bytes_avoided_fwd": 236 * (n_h + n_s) * 2}))


if __name__ == "__main__":
    triplane()
    stats()
    two_segments()
