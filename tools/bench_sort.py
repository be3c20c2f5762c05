"""Time the stable tile-bit sort (hugs_b200_sort_pairs_range on bits [32,45)) on pairs shaped like the
rasterizer's: n pairs emitted Gaussian by Gaussian, each a small rectangle of consecutive tile ids.
usage: python tools/bench_sort.py [n] [iters]   (HB_SORT_* env vars select kernel variants)"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ml-hugs_b200"))
import torch

from hugs_b200 import _native as N


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 12_250_000
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    g = torch.Generator(device="cuda").manual_seed(1)
    # rectangles of 3x4 tiles on a 120x68 grid, in "depth order" (random positions)
    ng = n // 12 + 1
    x0 = torch.randint(0, 117, (ng,), device="cuda", generator=g)
    y0 = torch.randint(0, 65, (ng,), device="cuda", generator=g)
    dx = torch.arange(12, device="cuda") % 3
    dy = torch.arange(12, device="cuda") // 3
    tile = ((y0[:, None] + dy[None]) * 120 + x0[:, None] + dx[None]).reshape(-1)[:n].long()
    depth = torch.arange(ng, device="cuda").repeat_interleave(12)[:n].long()
    keys = (tile << 32) | depth
    vals = torch.arange(n, device="cuda", dtype=torch.int32)
    tb = int(N.sort_temp_bytes(n))
    temp = torch.empty(tb, dtype=torch.uint8, device="cuda")
    k_in, v_in = keys.clone(), vals.clone()
    k_out, v_out = torch.empty_like(keys), torch.empty_like(vals)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def run():
        k_in.copy_(keys)
        v_in.copy_(vals)
        return N.sort_pairs_range(N.ptr(k_in), N.ptr(v_in), N.ptr(k_out), N.ptr(v_out), n, 32, 45,
                                  temp.data_ptr(), tb, st)

    def copies():
        k_in.copy_(keys)
        v_in.copy_(vals)

    for _ in range(3):
        assert run() == 0
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    for _ in range(iters):
        copies()
    e[1].record()
    for _ in range(iters):
        run()
    e[2].record()
    torch.cuda.synchronize()
    t_copy = e[0].elapsed_time(e[1]) / iters
    t_all = e[1].elapsed_time(e[2]) / iters
    ok = "skipped"
    if not os.environ.get("HB_SORT_NOLOOK"):
        order = torch.sort(tile, stable=True)[1]
        ok = bool((k_out == keys[order]).all() and (v_out == vals[order]).all())
    ms = t_all - t_copy
    print("n=%d sort[32,45) %.3f ms (hist + 2 passes; %.1f GB/s of 2x24 B/pair) correct=%s env=%s" % (
        n, ms, n * 48 / ms / 1e6, ok, {k: v for k, v in os.environ.items() if k.startswith("HB_SORT")}))


if __name__ == "__main__":
    main()
