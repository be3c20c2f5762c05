"""CUDA-graph capture of a whole data-parallel step (hugs_b200.dp.GraphedStep): the rasterizer's forward call is
capture-aware (csrc/api.cu: under capture it is enqueued for the remembered pair capacity and never waits for
the pair count), so forward + backward of several views over two streams replay as ONE graph.  The replayed
gradients equal the eager step's (summation order of the reductions aside)."""
import pytest
import torch

import refutil
import scenes

pytestmark = pytest.mark.gpu
NAMES = ("means3D", "shs", "opacity", "scales", "rotations")


def _setup(n_h, n_s, W, H, n_views, seed):
    from hugs_b200 import dp
    views = [scenes.make_scene(n_s, n_h, W, H, 300.0, seed=seed, device="cuda", view=v, n_views=n_views)
             for v in range(n_views)]
    params = {}
    for k in NAMES:
        params["h_" + k] = views[0][k][:n_h].detach().clone().contiguous().requires_grad_(True)
        params["s_" + k] = views[0][k][n_h:].detach().clone().contiguous().requires_grad_(True)
    buf = dp.FlatGradBuffer(params)
    buf.attach_sinks(list(params))
    return views, params, buf


def _render_view(views, params):
    import diff_gaussian_rasterization as dgr
    from hugs_b200 import segments

    def render(v):
        sc = views[v]
        seg = lambda t: tuple(params[t + k] for k in NAMES)  # noqa: E731
        P = params["h_means3D"].shape[0] + params["s_means3D"].shape[0]
        m2d = torch.empty(P, 3, device="cuda", requires_grad=True)
        color, _ = segments.rasterize_two_segments(seg("h_"), seg("s_"), m2d, scenes.settings_of(sc, dgr))
        return (color * sc["dL_dcolor"]).sum()
    return render


def test_graphed_step_replays_the_eager_gradients():
    from hugs_b200 import dp
    n_views = 4
    views, params, buf = _setup(3000, 7000, 320, 200, n_views, seed=71)
    render = _render_view(views, params)
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    ids = list(range(n_views))
    try:
        dp.dp_step(render, ids, buf, 0, 1, streams=streams, read_loss=False)
        torch.cuda.synchronize()
        ref = buf.flat.clone()
        assert float(ref.abs().max()) > 0
        step = dp.GraphedStep(lambda: dp.dp_step(render, ids, buf, 0, 1, streams=streams, read_loss=False,
                                                 exchange=False))
        for _ in range(5):
            losses = step.replay()
            torch.cuda.synchronize()
            err = refutil.rel_err(buf.flat.cpu().numpy(), ref.cpu().numpy())
            assert err < 1e-5, err
            assert len(losses) == n_views and all(bool(torch.isfinite(x)) for x in losses)
        # new inputs at the same addresses: the replay follows them
        for sc in views:
            sc["dL_dcolor"].mul_(2.0)
        step.replay()
        torch.cuda.synchronize()
        err = refutil.rel_err(buf.flat.cpu().numpy(), 2.0 * ref.cpu().numpy())
        assert err < 1e-5, err
    finally:
        import hugs_b200
        for t in params.values():
            hugs_b200.detach_grad_sink(t)


def test_capture_without_a_known_capacity_is_refused():
    """The first call for a (P, W, H) cannot be captured: no pair capacity is known yet (clear error, and the
    capture is left in a state torch can end)."""
    import diff_gaussian_rasterization as dgr
    sc = scenes.make_scene(1237, 0, 136, 72, 150.0, seed=72, device="cuda")   # a (P, W, H) no other test uses
    settings = scenes.settings_of(sc, dgr)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with pytest.raises(RuntimeError, match="capture"):
        with torch.cuda.graph(g):
            with torch.no_grad():
                dgr.GaussianRasterizer(settings)(
                    means3D=sc["means3D"], means2D=torch.zeros_like(sc["means3D"]), opacities=sc["opacity"],
                    shs=sc["shs"], scales=sc["scales"], rotations=sc["rotations"])
    torch.cuda.synchronize()


def test_graphed_step_without_sinks_over_two_streams():
    """Parameters without gradient sinks (autograd's AccumulateGrad) and views on two streams: dp_step pins the
    accumulation nodes to the main stream, which orders the additions on the device -- the captured step
    replays the eager gradients (without the pinning it replayed garbage within 50 steps)."""
    import diff_gaussian_rasterization as dgr
    from hugs_b200 import dp
    n_views = 4
    views = [scenes.make_scene(9000, 0, 320, 200, 300.0, seed=74, device="cuda", view=v, n_views=n_views)
             for v in range(n_views)]
    params = {k: views[0][k].detach().clone().contiguous().requires_grad_(True) for k in NAMES}
    buf = dp.FlatGradBuffer(params, gather=True)

    def render(v):
        sc = views[v]
        color, _ = dgr.GaussianRasterizer(scenes.settings_of(sc, dgr))(
            means3D=params["means3D"], means2D=torch.zeros_like(params["means3D"], requires_grad=True),
            opacities=params["opacity"], shs=params["shs"], scales=params["scales"], rotations=params["rotations"])
        return (color * sc["dL_dcolor"]).sum()

    streams = dp.view_streams(2, prioritise=False)
    ids = list(range(n_views))
    dp.dp_step(render, ids, buf, 0, 1, streams=streams, read_loss=False)
    torch.cuda.synchronize()
    ref = {k: p.grad.clone() for k, p in params.items()}
    step = dp.GraphedStep(lambda: dp.dp_step(render, ids, buf, 0, 1, streams=streams, read_loss=False, exchange=False))
    adopted = {k: p.grad for k, p in params.items()}   # the tensors autograd adopted as .grad during the capture
    for _ in range(30):
        step.replay()
    torch.cuda.synchronize()
    for k in NAMES:
        assert refutil.rel_err(adopted[k].cpu().numpy(), ref[k].cpu().numpy()) < 1e-5, k


@pytest.mark.parametrize("workload,two_seg", [("smoke_256", True), ("smoke_human_256", False)])
def test_bench_graphed_e2e_step_reads_the_same_losses_and_gradients_as_the_eager_loop(monkeypatch, workload, two_seg):
    """bench.py's end-to-end step with the graph (inputs prefetched into staging blocks, device-to-device copy,
    replay, losses read two steps later) against its eager e2e loop, same workload and inputs."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    dev = torch.device("cuda", 0)

    def run(graph):
        monkeypatch.setenv("HB_GRAPH", "1" if graph else "0")
        wl = bench.make_workload(workload, 0, 1, dev, views=4)
        rep = bench.Replica(wl, dev, bench.our_impl(), sh_sink=True)
        assert rep.two_seg == two_seg and rep.use_graph == graph
        rep.use_graph = False
        rep.step_resident()          # what bench.main does first: an eager step
        rep.use_graph = graph
        for _ in range(4):
            rep.step_e2e()
        rep.drain_e2e()
        torch.cuda.synchronize()
        grads = torch.cat([g.detach().reshape(-1) for g in rep.buf.grads().values() if g is not None]).clone()
        return list(rep.losses), grads

    l0, g0 = run(False)
    l1, g1 = run(True)
    assert len(l0) == len(l1) == 16
    for a, b in zip(l0, l1):
        assert abs(a - b) <= 1e-5 * max(abs(a), 1.0), (a, b)
    assert refutil.rel_err(g1.cpu().numpy(), g0.cpu().numpy()) < 1e-5


def test_blend_priority_changes_the_launch_and_nothing_else():
    """hugs_b200.dp.view_streams: high-priority streams + the blend kernels launched at the lowest priority
    (hugs_b200_set_blend_priority, cudaLaunchKernelEx with a priority attribute): same image bit for bit,
    same gradients up to the order of the reductions."""
    import diff_gaussian_rasterization as dgr
    from hugs_b200 import _native as N, dp
    sc = scenes.make_scene(6000, 2000, 320, 200, 300.0, seed=73, device="cuda")
    settings = scenes.settings_of(sc, dgr)

    def render():
        leaves = {k: sc[k].detach().clone().requires_grad_(True) for k in NAMES}
        color, _ = dgr.GaussianRasterizer(settings)(
            means3D=leaves["means3D"], means2D=torch.zeros_like(leaves["means3D"], requires_grad=True),
            opacities=leaves["opacity"], shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"])
        (color * sc["dL_dcolor"]).sum().backward()
        return color.detach(), leaves

    c0, l0 = render()
    torch.cuda.synchronize()
    streams = dp.view_streams(2)
    try:
        streams[0].wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(streams[0]):
            c1, l1 = render()
        torch.cuda.synchronize()
    finally:
        assert N.set_blend_priority(0, 0) == 1
    assert torch.equal(c0, c1)
    for k in NAMES:
        assert refutil.rel_err(l1[k].grad.cpu().numpy(), l0[k].grad.cpu().numpy()) < 1e-5, k
