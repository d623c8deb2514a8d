"""exg_colsum (bias gradients) on the gradient tensors of the TAN / ExoGround steps: us per launch and GB/s, graph-timed."""
import sys

import torch

sys.path.insert(0, ".")
from exoground_b200 import ops  # noqa: E402

K = ops.get()
print(f"{'M':>6} {'N':>6} {'us':>8} {'GB/s':>8}")
for M, N in [(8192, 1536), (8192, 2048), (12288, 1536), (12288, 2048), (16384, 1536), (17408, 2048), (12288, 512)]:
    x = torch.randn(M, N, device="cuda").bfloat16()
    out = torch.zeros(N, device="cuda")
    for _ in range(3):
        K.colsum(x, out=out)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        K.colsum(x, out=out)
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(10):
            K.colsum(x, out=out)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 50 * 1e3
    print(f"{M:>6} {N:>6} {us:>8.1f} {M * N * 2 / us / 1e3:>8.0f}", flush=True)
