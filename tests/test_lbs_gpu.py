"""Fused LBS kernel (forward + backward) against the golden fixture produced by the reference's own
Python chain, against that chain run live on the GPU (oracle/_ref/hugs_ref), and the batched
pose-frame path."""
import os

import numpy as np
import pytest
import torch

import refutil

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "lbs_golden.npz")


def rel(a, b):
    a = a.detach().cpu().double() if isinstance(a, torch.Tensor) else torch.from_numpy(np.asarray(a)).double()
    b = b.detach().cpu().double() if isinstance(b, torch.Tensor) else torch.from_numpy(np.asarray(b)).double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("tag", ["a", "b"])
def test_lbs_forward_backward_vs_reference_golden(tag):
    import hugs_b200

    g = np.load(GOLD)
    t = lambda k: torch.from_numpy(g[tag + "_" + k]).cuda()  # noqa: E731
    A, xyz, w, R, sc = [t(k).requires_grad_(True) for k in ("A", "xyz", "weights", "rotmat", "scales")]
    s = float(g[tag + "_smpl_scale"])
    oxyz, oq, osc = hugs_b200.deform_gaussians(A, xyz, w, R, sc, smpl_scale=s, transl=t("transl"))
    assert rel(oxyz, g[tag + "_out_xyz"]) < 1e-5
    assert float((oq.cpu() - torch.from_numpy(g[tag + "_out_rotq"])).abs().max()) < 1e-5
    assert rel(osc, g[tag + "_out_scales"]) < 1e-6
    loss = (oxyz * t("g_xyz")).sum() + (oq * t("g_rotq")).sum() + (osc * t("g_scales")).sum()
    loss.backward()
    for name, p in (("dA", A), ("d_xyz", xyz), ("d_weights", w), ("d_rotmat", R), ("d_scales", sc)):
        assert rel(p.grad, g[tag + "_" + name]) < 1e-3, name   # BASELINE: gradients 1e-3 relative


def test_matrix_to_quaternion_known_answers():
    import hugs_b200

    g = np.load(GOLD)
    q = hugs_b200.matrix_to_quaternion(torch.from_numpy(g["m2q_in"]).cuda())
    assert float((q.cpu() - torch.from_numpy(g["m2q_out"])).abs().max()) < 1e-6


def test_lbs_extra_signature_path_materialises_T():
    import hugs_b200

    g = np.load(GOLD)
    t = lambda k: torch.from_numpy(g["a_" + k]).cuda()  # noqa: E731
    A = t("A").requires_grad_(True)
    v = t("xyz").requires_grad_(True)
    w = t("weights").requires_grad_(True)
    verts, A_out, T, v_posed, v_shaped = hugs_b200.lbs_extra(A[None], v[None], None, w, torch.zeros(1, 72).cuda(),
                                                             disable_posedirs=True)
    assert verts.shape == (1, v.shape[0], 3) and T.shape == (1, v.shape[0], 4, 4)
    assert float((T[0].cpu() - torch.from_numpy(g["a_out_T"])).abs().max()) < 2e-6
    ref_verts = (torch.from_numpy(g["a_out_xyz"]) - torch.from_numpy(g["a_transl"])[None])
    assert float((verts[0].cpu() - ref_verts).abs().max()) < 1e-5
    (verts.sum() + (T * T).sum()).backward()
    assert A.grad is not None and v.grad is not None and w.grad is not None


@pytest.mark.skipif(not os.path.isdir(os.path.join(refutil.REF_DIR, "hugs_ref")), reason="reference LBS not installed")
def test_full_size_vs_reference_chain_live():
    """110 210 Gaussians (BASELINE configs[1]) against the reference's PyTorch chain on the same GPU."""
    import hugs_b200
    from hugs_b200 import synth

    lbs, rot = refutil.load_ref_lbs()
    h = synth.make_human_gaussians(synth.N_HUMAN, seed=77, device="cuda")
    leaves = lambda: [h[k].clone().requires_grad_(True) for k in ("xyz", "lbs_weights", "rotmat", "scales")] + \
        [h["A"][0].clone().requires_grad_(True)]  # noqa: E731
    gx = torch.randn(synth.N_HUMAN, 3, device="cuda")
    gq = torch.randn(synth.N_HUMAN, 4, device="cuda")
    gs = torch.randn(synth.N_HUMAN, 3, device="cuda")
    s = 1.3
    xyz, w, R, sc, A = leaves()
    oxyz, oq, osc = hugs_b200.deform_gaussians(A, xyz, w, R, sc, smpl_scale=s, transl=h["transl"][0])
    ((oxyz * gx).sum() + (oq * gq).sum() + (osc * gs).sum()).backward()
    xyz2, w2, R2, sc2, A2 = leaves()
    verts, _, T, _, _ = lbs.lbs_extra(A2[None], xyz2[None], None, w2, torch.zeros(1, 72).cuda(),
                                      disable_posedirs=True, pose2rot=True)
    rxyz = verts.squeeze(0) * s + h["transl"][0][None]
    rq = rot.matrix_to_quaternion(T.squeeze(0)[:, :3, :3] @ R2)
    rsc = sc2 * s
    ((rxyz * gx).sum() + (rq * gq).sum() + (rsc * gs).sum()).backward()
    assert rel(oxyz, rxyz) < 1e-5 and float((oq - rq).abs().max()) < 2e-5 and rel(osc, rsc) < 1e-6
    for a, b, name in ((xyz, xyz2, "xyz"), (w, w2, "weights"), (R, R2, "rotmat"), (sc, sc2, "scales"), (A, A2, "A")):
        assert rel(a.grad, b.grad) < 1e-3, name


def test_batched_pose_frames_equal_per_frame_calls():
    import hugs_b200
    from hugs_b200 import synth

    F = 5
    h = synth.make_human_gaussians(3001, seed=78, n_frames=F, device="cuda")
    bx, bq, bs = hugs_b200.lbs_deform_frames(h["A"], h["xyz"], h["lbs_weights"], h["rotmat"], h["scales"],
                                             h["transl"], 1.0)
    assert bx.shape == (F, 3001, 3) and bq.shape == (F, 3001, 4) and bs.shape == (F, 3001, 3)
    for f in range(F):
        x, q, s = hugs_b200.deform_gaussians(h["A"][f], h["xyz"], h["lbs_weights"], h["rotmat"], h["scales"],
                                             smpl_scale=1.0, transl=h["transl"][f])
        assert torch.equal(bx[f], x) and torch.equal(bq[f], q) and torch.equal(bs[f], s)
