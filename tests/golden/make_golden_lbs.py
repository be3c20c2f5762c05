#!/usr/bin/env python3
"""Generate tests/golden/lbs_golden.npz by running the REFERENCE's own Python on the CPU.

Run in the build container (needs /root/reference; it cannot travel to the GPU box, the fixture
does):   python tests/golden/make_golden_lbs.py

What runs is the reference chain for one pose, unmodified, loaded by file path:
  lbs_extra                (hugs/models/modules/lbs.py:19-73, disable_posedirs=True as in the release cfg)
  deformation tail         (hugs/models/hugs_trimlp.py:510-518, restated literally below)
  matrix_to_quaternion     (hugs/utils/rotations.py:105-156)
and torch.autograd through it for the gradient fixtures.  `smplx.lbs` (un-vendored dependency,
smplx==0.1.28) is shimmed by the reference's own vendored copy (hugs/datasets/neuman_utils/smpl.py:363-505).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "ml-hugs_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
REF = os.environ.get("HUGS_REFERENCE", "/root/reference")


def main():
    import refutil
    # synth is pure torch; import it without loading the CUDA library
    import importlib.util
    spec = importlib.util.spec_from_file_location("synth", os.path.join(ROOT, "ml-hugs_b200", "hugs_b200", "synth.py"))
    synth = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(synth)

    lbs, rot = refutil.load_ref_lbs(ref_root=REF)
    torch.manual_seed(0)
    out = {}
    for tag, N, seed, smpl_scale in (("a", 1536, 31, 1.0), ("b", 777, 32, 1.7)):
        h = synth.make_human_gaussians(N, seed=seed)
        A = h["A"][0].clone().requires_grad_(True)                    # [24,4,4]
        xyz = h["xyz"].clone().requires_grad_(True)
        w = h["lbs_weights"].clone().requires_grad_(True)
        R = h["rotmat"].clone().requires_grad_(True)
        sc = h["scales"].clone().requires_grad_(True)
        transl = h["transl"][0].clone()
        pose = torch.zeros(1, 72)
        # --- reference chain (hugs_trimlp.py:470-518) ---
        deformed_xyz, _, lbs_T, _, _ = lbs.lbs_extra(A[None], xyz[None], None, w, pose, disable_posedirs=True,
                                                     pose2rot=True)
        deformed_xyz = deformed_xyz.squeeze(0)
        lbs_T = lbs_T.squeeze(0)
        deformed_xyz = deformed_xyz * smpl_scale
        gs_scales = sc * smpl_scale
        deformed_xyz = deformed_xyz + transl[None]
        deformed_gs_rotmat = lbs_T[:, :3, :3] @ R
        deformed_gs_rotq = rot.matrix_to_quaternion(deformed_gs_rotmat)
        # --- seeded upstream gradients, autograd through the reference chain ---
        g = torch.Generator().manual_seed(seed + 1000)
        g_xyz = torch.randn(N, 3, generator=g)
        g_q = torch.randn(N, 4, generator=g)
        g_s = torch.randn(N, 3, generator=g)
        loss = (deformed_xyz * g_xyz).sum() + (deformed_gs_rotq * g_q).sum() + (gs_scales * g_s).sum()
        loss.backward()
        n = lambda t: t.detach().numpy().astype(np.float32)  # noqa: E731
        out.update({
            tag + "_A": n(A), tag + "_xyz": n(xyz), tag + "_weights": n(w), tag + "_rotmat": n(R),
            tag + "_scales": n(sc), tag + "_transl": n(transl), tag + "_smpl_scale": np.float32(smpl_scale),
            tag + "_out_xyz": n(deformed_xyz), tag + "_out_rotq": n(deformed_gs_rotq), tag + "_out_scales": n(gs_scales),
            tag + "_out_T": n(lbs_T),
            tag + "_g_xyz": n(g_xyz), tag + "_g_rotq": n(g_q), tag + "_g_scales": n(g_s),
            tag + "_dA": n(A.grad), tag + "_d_xyz": n(xyz.grad), tag + "_d_weights": n(w.grad),
            tag + "_d_rotmat": n(R.grad), tag + "_d_scales": n(sc.grad),
        })
    # matrix_to_quaternion known answers incl. the degenerate / clamped branches
    mats = torch.stack([
        torch.eye(3), torch.diag(torch.tensor([1.0, -1.0, -1.0])), torch.diag(torch.tensor([-1.0, 1.0, -1.0])),
        torch.diag(torch.tensor([-1.0, -1.0, 1.0])), 0.05 * torch.eye(3), torch.zeros(3, 3),
        torch.tensor([[0.0, -1.0, 0.0], [1.0, 0.0, 0.0], [0.0, 0.0, 1.0]]),
        0.8 * torch.tensor([[0.36, 0.48, -0.8], [-0.8, 0.6, 0.0], [0.48, 0.64, 0.6]]),
    ])
    out["m2q_in"] = mats.numpy()
    out["m2q_out"] = rot.matrix_to_quaternion(mats).numpy()
    path = os.path.join(HERE, "lbs_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
