"""GPU-box timing of the fused L1+SSIM loss (forward + backward at 1920x1080) against the plain-torch
evaluation the reference runs (hugs/losses/utils.py: five depthwise conv2d + elementwise + autograd).
Prints one JSON line; algorithmic bytes: forward 8 B read + 12 B written, backward 20 B read + 4 B
written per pixel and channel (44 B x 3 x H x W)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "ml-hugs_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch

from hugs_b200 import losses
from test_loss_gpu import _torch_ssim


def timeit(fn, iters=30, warm=5):
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


def main():
    H, W = 1080, 1920
    gt = torch.rand(3, H, W, device="cuda")
    pred = (gt + 0.2 * (torch.rand(3, H, W, device="cuda") - 0.5)).clamp(0, 1).requires_grad_(True)

    def ours():
        pred.grad = None
        l1, s = losses.l1_ssim(pred, gt)
        (0.8 * l1 + 0.2 * (1.0 - s)).backward()

    def ref():
        pred.grad = None
        l1 = (pred - gt).abs().mean()
        s = _torch_ssim(pred[None], gt[None])
        (0.8 * l1 + 0.2 * (1.0 - s)).backward()

    t_ours, t_ref = timeit(ours), timeit(ref)
    peak = 6551.4
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    nbytes = 44 * 3 * H * W
    print(json.dumps({"op": "l1+ssim fwd+bwd 3x1080x1920", "ours_ms": round(t_ours, 4), "torch_reference_ms": round(t_ref, 4),
                      "speedup": round(t_ref / t_ours, 2), "algorithmic_bytes": nbytes,
                      "gbs": round(nbytes / t_ours / 1e6, 1), "hbm_frac": round(nbytes / t_ours / 1e6 / peak, 4)}))


if __name__ == "__main__":
    main()
