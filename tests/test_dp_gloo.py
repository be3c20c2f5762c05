"""World-size-2 gloo tests of the view-batch data-parallel host logic (hugs_b200/dp.py): view
sharding, the flat gradient buffer, and that the allreduced gradient equals the single-process
gradient of the summed loss.  The per-view loss here is a small differentiable stand-in (the CUDA
rasterizer cannot run on the CPU box); the GPU box exercises the real path in bench.py --gpus N."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _params(seed=3):
    g = torch.Generator().manual_seed(seed)
    return {"means": torch.randn(50, 3, generator=g).requires_grad_(True),
            "scales": torch.rand(50, 3, generator=g).requires_grad_(True),
            "shs": torch.randn(50, 16, 3, generator=g).requires_grad_(True)}


def _loss(params, view):
    c = torch.cos(torch.tensor(0.3 * view))
    proj = params["means"][:, :2] * c / (params["means"][:, 2:3].abs() + 1.0)
    return (proj.sum() + (params["scales"] * (view + 1)).pow(2).sum() + (params["shs"][:, 0] * c).sum()
            + (params["shs"][:, 1 + view % 15] ** 2).sum())


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_views, out, gather=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "ml-hugs_b200"))
    from hugs_b200 import dp

    r, _, w = dp.init_from_env("gloo")
    assert (r, w) == (rank, world)
    params = _params()
    buf = dp.FlatGradBuffer(params, gather=gather)
    views = list(range(n_views))
    for _ in range(2):  # two steps: the buffer must be cleared between steps
        local = dp.dp_step(lambda v: _loss(params, v), views, buf, rank, world)
    tot = dp.sum_over_ranks([local])[0]
    mx = dp.max_over_ranks(float(rank))
    if rank == 0:
        torch.save({"grads": {k: v.clone() for k, v in buf.grads().items()}, "loss": tot, "max": mx}, out)
    dist.destroy_process_group()


@pytest.mark.parametrize("gather", [False, True])
@pytest.mark.parametrize("n_views", [4, 5])
def test_dp_allreduce_equals_single_process_gradient(tmp_path, n_views, gather):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), n_views, out, gather), nprocs=2, join=True)
    got = torch.load(out)
    params = _params()
    loss = sum(_loss(params, v) for v in range(n_views))
    loss.backward()
    for k, p in params.items():
        assert torch.allclose(got["grads"][k], p.grad, rtol=1e-5, atol=1e-6), k
    assert abs(got["loss"] - float(loss)) < 1e-3 * abs(float(loss))
    assert got["max"] == 1.0


def test_shard_views_partition():
    from hugs_b200 import dp

    for world in (1, 2, 4, 8):
        for n in (0, 1, 7, 8, 300):
            shards = [dp.shard_views(n, r, world) for r in range(world)]
            flat = sorted(v for s in shards for v in s)
            assert flat == list(range(n))
            assert max(len(s) for s in shards) - min(len(s) for s in shards) <= 1
    with pytest.raises(ValueError):
        dp.shard_views(4, 2, 2)


def test_flat_grad_buffer_views_and_alignment():
    from hugs_b200 import dp

    params = _params()
    buf = dp.FlatGradBuffer(params)
    assert all(o % 32 == 0 for o in buf.offsets) and buf.nbytes >= 4 * sum(p.numel() for p in params.values())
    _loss(params, 1).backward()
    ref = _params()
    _loss(ref, 1).backward()
    for k in params:
        assert params[k].grad.data_ptr() >= buf.flat.data_ptr()
        assert torch.allclose(params[k].grad, ref[k].grad)
    assert buf.allreduce() is None  # no process group: no-op
    buf.zero_()
    assert float(buf.flat.abs().sum()) == 0.0
    with pytest.raises(ValueError):
        dp.FlatGradBuffer({"x": torch.zeros(3, dtype=torch.float64, requires_grad=True)})


def _view_stats(view, n=40):
    g = torch.Generator().manual_seed(100 + view)
    radii = torch.randint(-3, 30, (n,), generator=g, dtype=torch.int32).clamp_min(0)
    grad = torch.randn(n, 3, generator=g)
    return radii, grad


def _apply_reference_lines(radii, grad, acc, den, mx):
    vis = radii > 0                                     # gs_renderer.py:156
    mx[vis] = torch.max(mx[vis], radii[vis].float())    # gs_trainer.py:406-411
    acc[vis] += torch.norm(grad[vis, :2], dim=-1, keepdim=True)   # scene.py:460-461
    den[vis] += 1                                       # scene.py:462


def _stats_worker(rank, world, port, n_views, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "ml-hugs_b200"))
    from hugs_b200 import dp

    dp.init_from_env("gloo")
    n = 40
    acc, den, mx = torch.full((n, 1), 0.5), torch.ones(n, 1), torch.full((n,), 2.0)
    sync = dp.DensificationStatsSync(n)
    for step in range(2):
        sync.begin_step()
        for v in dp.shard_views(n_views, rank, world):
            radii, grad = _view_stats(10 * step + v)
            _apply_reference_lines(radii, grad, sync.step_accum, sync.step_denom, mx)   # CPU stand-in of add_view
        sync.finish_step(acc, den, mx)
    torch.save({"acc": acc, "den": den, "mx": mx}, out + str(rank))
    dist.destroy_process_group()


@pytest.mark.parametrize("n_views", [4, 5])
def test_densification_stats_sync_equals_single_process(tmp_path, n_views):
    out = str(tmp_path / "stats")
    mp.spawn(_stats_worker, args=(2, _free_port(), n_views, out), nprocs=2, join=True)
    n = 40
    acc, den, mx = torch.full((n, 1), 0.5), torch.ones(n, 1), torch.full((n,), 2.0)
    for step in range(2):
        for v in range(n_views):
            _apply_reference_lines(*_view_stats(10 * step + v), acc, den, mx)
    for rank in range(2):
        got = torch.load(out + str(rank))
        assert torch.equal(got["den"], den) and torch.equal(got["mx"], mx)
        assert torch.allclose(got["acc"], acc, rtol=1e-6, atol=1e-6)
