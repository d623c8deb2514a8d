#!/usr/bin/env python
"""Times exg_nce_logdiff + exg_tan_loss_tail at the bench shape (B=128, T=64, N=32, E=D=6)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from exoground_b200 import ops  # noqa: E402

K = ops.get()
B, T, N, E, D = 128, 64, 32, 6, 6
dev = "cuda"
g = torch.Generator().manual_seed(0)
tgt = (torch.rand((B, N, T), generator=g) < 0.15).to(torch.uint8).to(dev)
cv = (torch.rand((B * N,), generator=g) < 0.8).to(torch.uint8).to(dev)
tgt = (tgt * cv.view(B, N, 1)).contiguous()


def sums(S):
    rows = torch.rand((2, S, B * T), generator=g).to(dev) + 0.5
    cols = torch.rand((2, S, B * N), generator=g).to(dev) + 0.5
    return rows[0], rows[1], cols[0], cols[1]


sd, sj = sums(E), sums(D)
md, mj = torch.randn(B * N, generator=g).to(dev), torch.randn(B * N, generator=g).to(dev)
la = torch.randn(B * N, generator=g).to(dev)
for th, head in ((0.5, 1), (0.0, 1), (0.5, 0)):
    for _ in range(3):
        K.tan_loss_tail(tgt, cv, sd, sj, md, mj, la if head else None, None, B, T, N, 1 / 0.07, th, head, 1.0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        K.tan_loss_tail(tgt, cv, sd, sj, md, mj, la if head else None, None, B, T, N, 1 / 0.07, th, head, 1.0)
    e1.record()
    torch.cuda.synchronize()
    print(f"threshold={th} head={head}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us per call (2 launches + allocations)")
