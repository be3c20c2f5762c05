"""Seeded test scenes and the three runners (our CUDA path, the reference build, the CPU oracle)
that the parity tests compare.  Test infrastructure only."""
import numpy as np
import torch

from hugs_b200 import synth


def make_scene(P_scene, P_human, W, H, focal, seed, degree=3, device="cpu", view=0, n_views=1):
    """human || scene concatenation as hugs/renderer/gs_renderer.py:33-37 (human first)."""
    cam = synth.make_camera(W, H, focal, view=view, n_views=n_views, seed=seed)
    parts = []
    if P_human > 0:
        h = synth.make_human_gaussians(P_human, seed + 1)
        # deform on the CPU with plain torch (only to place the human in front of the camera)
        T = torch.einsum("nj,jab->nab", h["lbs_weights"], h["A"][0])
        xyz = (T[:, :3, :3] @ h["xyz"][..., None])[..., 0] + T[:, :3, 3] + h["transl"][0]
        Rp = T[:, :3, :3] @ h["rotmat"]
        q = _mat2quat(Rp)
        parts.append(dict(means3D=xyz, scales=h["scales"], rotations=q, opacity=h["opacity"], shs=h["shs"]))
    if P_scene > 0:
        parts.append(synth.make_scene_gaussians(P_scene, seed + 2, cam["tanfovx"], cam["tanfovy"]))
    keys = ["means3D", "scales", "rotations", "opacity", "shs"]
    sc = {k: torch.cat([p[k] for p in parts], dim=0).contiguous() for k in keys}
    sc["dL_dcolor"] = synth.make_dL_dcolor(W, H, seed)
    sc["degree"] = degree
    sc.update(cam)
    dev = torch.device(device)
    return {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}


def _mat2quat(m):
    """plain-torch matrix_to_quaternion (hugs/utils/rotations.py:105-156), used only to build inputs."""
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = [m[:, i, j] for i in range(3) for j in range(3)]
    x = torch.stack([1 + m00 + m11 + m22, 1 + m00 - m11 - m22, 1 - m00 + m11 - m22, 1 - m00 - m11 + m22], -1)
    qa = torch.where(x > 0, torch.sqrt(x.clamp_min(1e-30)), torch.zeros_like(x))
    cand = torch.stack([
        torch.stack([qa[:, 0] ** 2, m21 - m12, m02 - m20, m10 - m01], -1),
        torch.stack([m21 - m12, qa[:, 1] ** 2, m10 + m01, m02 + m20], -1),
        torch.stack([m02 - m20, m10 + m01, qa[:, 2] ** 2, m12 + m21], -1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, qa[:, 3] ** 2], -1)], -2)
    cand = cand / (2.0 * qa[..., None].clamp_min(0.1))
    best = qa.argmax(-1)
    return cand[torch.arange(m.shape[0]), best]


def settings_of(sc, module, debug=False):
    return module.GaussianRasterizationSettings(
        image_height=sc["image_height"], image_width=sc["image_width"], tanfovx=sc["tanfovx"],
        tanfovy=sc["tanfovy"], bg=sc["bg"], scale_modifier=1.0, viewmatrix=sc["viewmatrix"],
        projmatrix=sc["projmatrix"], sh_degree=sc["degree"], campos=sc["campos"], prefiltered=False,
        debug=debug)


def run_module(sc, module, backward=True, colors_precomp=None, cov3D_precomp=None):
    """Run a diff_gaussian_rasterization-shaped module (ours or the reference) through its public API
    and return outputs + gradients + the three opaque buffers."""
    names = ["means3D", "scales", "rotations", "opacity", "shs"]
    # inputs always require grad so that the autograd node (which holds the three opaque buffers and
    # num_rendered) exists even when backward is not run
    t = {k: sc[k].detach().clone().float().requires_grad_(True) for k in names}
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    settings = settings_of(sc, module)
    rast = module.GaussianRasterizer(settings)
    kw = dict(means3D=t["means3D"], means2D=means2D, opacities=t["opacity"])
    if colors_precomp is not None:
        cp = colors_precomp.detach().clone().requires_grad_(True)
        kw["colors_precomp"] = cp
    else:
        kw["shs"] = t["shs"]
    if cov3D_precomp is not None:
        c3 = cov3D_precomp.detach().clone().requires_grad_(True)
        kw["cov3D_precomp"] = c3
    else:
        kw["scales"] = t["scales"]
        kw["rotations"] = t["rotations"]
    color, radii = rast(**kw)
    out = dict(color=color.detach(), radii=radii.detach())
    # the autograd node keeps the buffers; fetch them for decoding
    node = color.grad_fn
    saved = node.saved_tensors if node is not None else None
    if saved is not None:
        out["geom"], out["binning"], out["img"] = saved[7], saved[8], saved[9]
        out["num_rendered"] = node.num_rendered if hasattr(node, "num_rendered") else None
    if backward:
        (color * sc["dL_dcolor"]).sum().backward()
        out["grads"] = dict(means3D=t["means3D"].grad, means2D=means2D.grad, opacity=t["opacity"].grad)
        if colors_precomp is None:
            out["grads"]["shs"] = t["shs"].grad
        else:
            out["grads"]["colors_precomp"] = cp.grad
        if cov3D_precomp is None:
            out["grads"]["scales"] = t["scales"].grad
            out["grads"]["rotations"] = t["rotations"].grad
        else:
            out["grads"]["cov3D_precomp"] = c3.grad
    return out


def run_oracle(sc, backward=True):
    from oracle import oracle as O

    n = lambda v: v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v  # noqa: E731
    W, H = sc["image_width"], sc["image_height"]
    kw = dict(shs=n(sc["shs"]), degree=sc["degree"], scales=n(sc["scales"]), rotations=n(sc["rotations"]))
    fw = O.forward(n(sc["means3D"]), n(sc["opacity"]), n(sc["viewmatrix"]), n(sc["projmatrix"]), n(sc["campos"]),
                   n(sc["bg"]), W, H, sc["tanfovx"], sc["tanfovy"], **kw)
    out = dict(fw=fw)
    if backward:
        out["grads"] = O.backward(fw, n(sc["dL_dcolor"]), n(sc["means3D"]), n(sc["viewmatrix"]), n(sc["projmatrix"]),
                                  n(sc["campos"]), n(sc["bg"]), W, H, sc["tanfovx"], sc["tanfovy"], **kw)
    return out
