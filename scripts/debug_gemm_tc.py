"""GPU debug helper: where do tcgen05 GEMM results differ from torch?  (not a test)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from exoground_b200 import ops
from exoground_b200.ops import lib as L

K = ops.get()
torch.manual_seed(0)

def case(M, N, Kd, tA, tB, reps=3):
    A = torch.randn((Kd, M) if tA else (M, Kd), device="cuda").bfloat16()
    B = torch.randn((Kd, N) if tB else (N, Kd), device="cuda").bfloat16()
    a = A.float().t() if tA else A.float()
    b = B.float() if tB else B.float().t()
    ref = a @ b
    for r in range(reps):
        got = K.gemm(A, B, transA=tA, transB=tB, impl=L.GEMM_TC, out_dtype=torch.float32)
        torch.cuda.synchronize()
        err = (got - ref).abs()
        bad = err > 0.05 * Kd ** 0.5
        msg = f"M{M} N{N} K{Kd} tA{int(tA)} tB{int(tB)} rep{r}: max err {err.max().item():.3f} bad frac {bad.float().mean().item():.4f}"
        if bad.any():
            rows = bad.any(1).nonzero().flatten()
            cols = bad.any(0).nonzero().flatten()
            msg += f" rows[{rows.min().item()}..{rows.max().item()}] n={rows.numel()} cols[{cols.min().item()}..{cols.max().item()}] n={cols.numel()}"
            # tile map
            tm = bad.view(-1, 1, bad.shape[1])[:(M // 128) * 128].view(M // 128, 128, N).any(1) if M % 128 == 0 else None
            # ratio got/ref at a bad location
            i = bad.nonzero()[0]
            msg += f" first bad {tuple(i.tolist())}: got {got[i[0], i[1]].item():.3f} ref {ref[i[0], i[1]].item():.3f}"
        print(msg, flush=True)

for shape in [(128, 128, 64), (256, 256, 512), (128, 128, 2048), (384, 512, 2048), (128, 256, 512), (128, 256, 64),
              (8192, 1536, 512), (8192, 256, 64), (19072, 256, 64), (19072, 128, 64)]:
    for tA, tB in [(False, False), (False, True), (True, True), (True, False)]:
        try:
            case(*shape, tA, tB)
        except Exception as e:
            print("EXC", shape, tA, tB, e)
