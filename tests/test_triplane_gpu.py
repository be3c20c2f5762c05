"""Tri-plane sampling kernels through the C ABI: against the golden vectors of the
reference's own module, against the CPU oracle, and at the reference's default size (32 features, 256^3,
110 210 points) against torch's F.grid_sample on the GPU.  Floating point: features within 1e-5 absolute
of torch's CUDA sampler (bit-exact against the CPU golden), gradients within 1e-4 relative to the largest
element (atomic accumulation order differs)."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import refutil  # noqa: F401
from oracle import triplane_oracle as TO

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "triplane_golden.npz")


def _module(planes, check_range=True):
    from hugs_b200.triplane import TriPlane
    pxy = torch.from_numpy(planes[0])
    tp = TriPlane(features=pxy.shape[1], resX=pxy.shape[2], resY=pxy.shape[3], resZ=planes[1].shape[3],
                  check_range=check_range)
    tp.load_state_dict({"plane_xy": pxy, "plane_xz": torch.from_numpy(planes[1]),
                        "plane_yz": torch.from_numpy(planes[2])})
    return tp.cuda()


@pytest.mark.parametrize("tag", ["a", "b", "c", "d"])
def test_matches_reference_golden(tag):
    g = np.load(GOLD)
    tp = _module([g[tag + k] for k in ("_pxy", "_pxz", "_pyz")])
    for p in tp.parameters():
        assert p.is_contiguous(memory_format=torch.channels_last) or p.shape[1] == 1
    x = torch.from_numpy(g[tag + "_x"]).cuda().requires_grad_(True)
    y = tp(x)
    assert np.array_equal(y.detach().cpu().numpy(), g[tag + "_feat"])        # bit-exact
    (y * torch.from_numpy(g[tag + "_w"]).cuda()).sum().backward()
    ref = g[tag + "_gx"]
    assert np.abs(x.grad.cpu().numpy() - ref).max() <= 1e-4 * np.abs(ref).max()
    for p, key in ((tp.plane_xy, "_gpxy"), (tp.plane_xz, "_gpxz"), (tp.plane_yz, "_gpyz")):
        ref = g[tag + key]
        assert p.grad.shape == ref.shape
        assert np.abs(p.grad.cpu().numpy() - ref).max() <= 1e-4 * max(np.abs(ref).max(), 1.0)


def test_batched_input_shape_and_plain_layout_planes():
    from hugs_b200.triplane import sample_triplane
    g = np.load(GOLD)
    planes = [torch.from_numpy(g["batched_" + k]).cuda() for k in ("pxy", "pxz", "pyz")]      # NCHW-contiguous
    y = sample_triplane(torch.from_numpy(g["batched_x"]).cuda(), *planes)
    assert y.shape == (2, 7, 12)
    assert np.array_equal(y.cpu().numpy(), g["batched_feat"])


def test_random_against_oracle_with_points_outside():
    rng = np.random.default_rng(3)
    planes = [rng.standard_normal(s).astype(np.float32) for s in ((1, 8, 7, 9), (1, 8, 7, 5), (1, 8, 9, 5))]
    x = (rng.random((777, 3), dtype=np.float32) * 3 - 1.5).astype(np.float32)      # a third lies outside
    tp = _module(planes, check_range=False)
    xt = torch.from_numpy(x).cuda().requires_grad_(True)
    y = tp(xt)
    assert np.array_equal(y.detach().cpu().numpy(), TO.forward(x, *planes))
    w = rng.standard_normal((777, 24)).astype(np.float32)
    (y * torch.from_numpy(w).cuda()).sum().backward()
    gx, gxy, gxz, gyz = TO.backward(x, *planes, w)
    assert np.abs(xt.grad.cpu().numpy() - gx).max() <= 1e-4 * np.abs(gx).max()
    for p, ref in ((tp.plane_xy, gxy), (tp.plane_xz, gxz), (tp.plane_yz, gyz)):
        assert np.abs(p.grad.cpu().numpy() - ref).max() <= 1e-4 * np.abs(ref).max()


def test_range_assertion_like_the_reference():
    g = np.load(GOLD)
    tp = _module([g["a" + k] for k in ("_pxy", "_pxz", "_pyz")])
    bad = torch.tensor([[0.0, 0.0, 0.0], [1.01, 0.0, 0.0]], device="cuda")
    with pytest.raises(AssertionError, match=r"x must be in \[0, 1\]"):
        tp(bad)
    tp.check_range = False
    assert tp(bad).shape == (2, 12)
    assert tp(torch.zeros(0, 3, device="cuda")).shape == (0, 12)
    with pytest.raises(RuntimeError, match="no CPU path"):
        tp(torch.zeros(2, 3))


def test_default_size_against_torch_grid_sample():
    """32 features, 256^3 planes, 110 210 points (the human model's size): values and all gradients
    against the reference's expression evaluated by torch on the same GPU."""
    from hugs_b200.triplane import TriPlane
    torch.manual_seed(7)
    tp = TriPlane().cuda()
    n = 110210
    x = (torch.rand(n, 3, device="cuda") * 2 - 1)
    x[:4] = torch.tensor([[-1.0, -1, -1], [1, 1, 1], [1, -1, 0], [0, 0, 0]], device="cuda")
    w = torch.randn(n, 96, device="cuda")
    xa = x.clone().requires_grad_(True)
    ya = tp(xa)
    (ya * w).sum().backward()
    ours = [xa.grad] + [p.grad.clone() for p in tp.parameters()]
    planes = [p.detach().clone().contiguous().requires_grad_(True) for p in tp.parameters()]
    xb = x.clone().requires_grad_(True)
    coords = (((xb - 0.0) / 2.0 + 0.5) * 2 - 1).reshape(1, -1, 1, 3)
    feats = [F.grid_sample(pl, coords[..., idx], align_corners=True)[0, :, :, 0].transpose(0, 1)
             for pl, idx in zip(planes, ([0, 1], [0, 2], [1, 2]))]
    yb = torch.cat(feats, dim=1)
    (yb * w).sum().backward()
    assert float((ya - yb).abs().max()) <= 1e-5
    ref = [xb.grad] + [p.grad for p in planes]
    for a, b in zip(ours, ref):
        assert a.shape == b.shape
        assert float((a - b).abs().max()) <= 1e-4 * float(b.abs().max())
