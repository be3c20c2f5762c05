"""Per-view duplicate / sort stage times and pair counts on the bench's orbit (diagnostic)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "ml-hugs_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch

import diff_gaussian_rasterization as dgr
import scenes
from hugs_b200 import _native as N

n_views = int(sys.argv[1]) if len(sys.argv) > 1 else 8
for v in range(n_views):
    sc = scenes.make_scene(1000000, 110210, 1920, 1080, 1400.0, seed=1002, device="cuda", view=v, n_views=n_views)
    settings = scenes.settings_of(sc, dgr)
    rast = dgr.GaussianRasterizer(settings)
    m2d = torch.zeros_like(sc["means3D"])
    kw = dict(means3D=sc["means3D"], means2D=m2d, shs=sc["shs"], opacities=sc["opacity"], scales=sc["scales"],
              rotations=sc["rotations"])
    with torch.no_grad():
        for _ in range(3):
            rast(**kw)
        torch.cuda.synchronize()
        N.stage_timing_enable(1)
        N.read_stage_times(reset=True)
        for _ in range(10):
            rast(**kw)
        torch.cuda.synchronize()
        t = N.read_stage_times(reset=True)
        N.stage_timing_enable(0)
    radii = rast(**kw)[1]
    print(v, {k: round(ms / max(c, 1), 4) for k, (ms, c) in t.items() if k in ("duplicate_with_keys", "sort_pairs", "render_forward")},
          "visible", int((radii > 0).sum()))
