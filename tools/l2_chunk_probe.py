"""Does chunk-major order keep a read + read-modify-write working set in the 126 MB L2?  (feasibility probe for
running the scene segment's per-Gaussian backward of all views of a step chunk by chunk)"""
import torch
P, V = 1_000_000, 8
sh = torch.randn(P, 48, device="cuda")
g = torch.zeros(P, 48, device="cuda")


def run(chunk):
    for c0 in range(0, P, chunk):
        a, b = g[c0:c0 + chunk], sh[c0:c0 + chunk]
        for _ in range(V):
            torch.add(a, b, out=a)


for chunk in (P, 400_000, 200_000, 131_072, 65_536, 32_768):
    for _ in range(2):
        run(chunk)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        run(chunk)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("chunk %8d rows (%5.1f MB working set): %.3f ms for %d passes, %.2f TB/s effective" % (
        chunk, chunk * 48 * 4 * 2 / 1e6, ms, V, V * P * 48 * 4 * 3 / ms / 1e9))
