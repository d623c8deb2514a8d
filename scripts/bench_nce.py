"""Times exg_nce_fwd / exg_nce_bwd at the bench shape (S=6, B=128, T=64, N=32, bf16) with CUDA events.
EXG_NO_NCE_TMA=1 selects the generic kernels."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from exoground_b200 import ops  # noqa: E402

K = ops.get()
S, B, T, N = 6, 128, 64, 32
dev = "cuda"
g = torch.Generator().manual_seed(0)
Lg = (torch.rand((S, B * T, B * N), generator=g) * 2 - 1).to(torch.bfloat16).to(dev)
tgt = (torch.rand((B, N, T), generator=g) < 0.1).to(torch.uint8).to(dev)
cv = (torch.rand((B * N,), generator=g) < 0.8).to(torch.uint8).to(dev)
tgt = tgt * cv.view(B, N, 1)
wr = torch.rand(B * T, generator=g).to(dev)
wc = torch.rand(B * N, generator=g).to(dev)
it = 1 / 0.07


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


bits = K.nce_pack_targets(tgt)
sums = K.nce_fwd(Lg, tgt, cv, B, T, N, it, tgt_bits=bits)
tf = timeit(lambda: K.nce_fwd(Lg, tgt, cv, B, T, N, it, tgt_bits=bits))
tb = timeit(lambda: K.nce_bwd(Lg, tgt, cv, sums, wr, wc, B, T, N, it, tgt_bits=bits))
nbytes = Lg.numel() * 2
print(f"nce_fwd {tf*1e3:8.1f} us  {nbytes/tf/1e6:8.1f} GB/s   (incl. 4 memsets)")
print(f"nce_bwd {tb*1e3:8.1f} us  {2*nbytes/tb/1e6:8.1f} GB/s")
print("checksums", [float(x.double().sum()) for x in sums])
