"""Fused L1 + SSIM kernels through the C ABI: against the golden vectors of the
reference's own Python, against the CPU oracle, and at 1080p against a plain torch fp32 restatement
of the same op on the GPU.  Tolerances (floating point): loss values 1e-5 absolute, gradients 1e-3
relative to the largest gradient element."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import refutil  # noqa: F401
from oracle import loss_oracle as LO

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "loss_golden.npz")


def _ours(pred, gt, w_l1=0.8, w_ssim=0.2):
    from hugs_b200 import losses
    p = torch.from_numpy(np.ascontiguousarray(pred)).cuda().requires_grad_(True) if not torch.is_tensor(pred) else pred
    g = torch.from_numpy(np.ascontiguousarray(gt)).cuda() if not torch.is_tensor(gt) else gt
    l1, s = losses.l1_ssim(p, g)
    (w_l1 * l1 + w_ssim * (1.0 - s)).backward()
    return float(l1), float(s), p.grad


@pytest.mark.parametrize("tag", ["a", "b", "c", "d", "e"])
def test_fused_loss_matches_reference_golden(tag):
    g = np.load(GOLD)
    l1, s, grad = _ours(g[tag + "_pred"], g[tag + "_gt"])
    assert abs(l1 - float(g[tag + "_l1"])) <= 1e-5
    assert abs(s - float(g[tag + "_ssim"])) <= 1e-5
    ref = g[tag + "_grad"]
    assert np.abs(grad.cpu().numpy() - ref).max() <= 1e-3 * np.abs(ref).max()


def test_fused_loss_matches_oracle_ragged_and_batched():
    rng = np.random.default_rng(5)
    for shape in [(3, 33, 17), (2, 3, 20, 45), (1, 1, 1), (3, 11, 11)]:
        gt = rng.random(shape, dtype=np.float32)
        pred = np.clip(gt + 0.3 * (rng.random(shape, dtype=np.float32) - 0.5), 0, 1).astype(np.float32)
        l1, s, grad = _ours(pred, gt)
        o = LO.l1_ssim(pred, gt, g_l1=0.8, g_ssim=-0.2)
        assert abs(l1 - o["l1"]) <= 1e-5 and abs(s - o["ssim"]) <= 1e-5
        assert np.abs(grad.cpu().numpy() - o["grad"]).max() <= 1e-3 * np.abs(o["grad"]).max()


def _torch_ssim(img1, img2):
    """plain torch fp32 restatement of hugs/losses/utils.py:63-106 (reference of the same op on the GPU)."""
    w1 = torch.from_numpy(LO.window_1d()).to(img1.device)
    C = img1.shape[-3]
    win = (w1[:, None] @ w1[None, :]).float()[None, None].expand(C, 1, 11, 11).contiguous()
    conv = lambda t: F.conv2d(t, win, padding=5, groups=C)  # noqa: E731
    mu1, mu2 = conv(img1), conv(img2)
    s1 = conv(img1 * img1) - mu1 * mu1
    s2 = conv(img2 * img2) - mu2 * mu2
    s12 = conv(img1 * img2) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))).mean()


def test_fused_loss_full_size_1080p_vs_torch():
    g = torch.Generator("cuda").manual_seed(3)
    gt = torch.rand(3, 1080, 1920, device="cuda", generator=g)
    pred = (gt + 0.2 * (torch.rand(3, 1080, 1920, device="cuda", generator=g) - 0.5)).clamp(0, 1)
    a = pred.clone().requires_grad_(True)
    l1, s, grad = _ours(a, gt)
    b = pred.clone().requires_grad_(True)
    rl1 = (b - gt).abs().mean()
    rs = _torch_ssim(b[None], gt[None])
    (0.8 * rl1 + 0.2 * (1.0 - rs)).backward()
    assert abs(l1 - float(rl1)) <= 1e-5 and abs(s - float(rs)) <= 1e-5
    assert float((grad - b.grad).abs().max()) <= 1e-3 * float(b.grad.abs().max())


def test_reference_named_wrappers_and_no_cpu_path():
    from hugs_b200 import losses
    gt = torch.rand(3, 40, 56, device="cuda")
    pred = torch.rand(3, 40, 56, device="cuda", requires_grad=True)
    l1, s = losses.l1_ssim(pred, gt)
    assert abs(float(losses.l1_loss(pred, gt)) - float(l1)) < 1e-7
    assert abs(float(losses.ssim(pred, gt)) - float(s)) < 1e-7
    mask = (torch.rand(40, 56, device="cuda") > 0.5).float()
    assert abs(float(losses.l1_loss(pred, gt, mask)) - float((pred - gt).abs().sum() / mask.sum())) < 1e-4
    with pytest.raises(NotImplementedError):
        losses.ssim(pred, gt, window_size=7)
    with pytest.raises(RuntimeError):
        losses.l1_ssim(pred.detach().cpu(), gt.cpu())
