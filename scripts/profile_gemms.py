"""Per-shape timing of every exg_gemm call of one eager bench step (CUDA events around each call)."""
import collections
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from exoground_b200 import ops  # noqa: E402
from exoground_b200.model import TemporalAligner  # noqa: E402
from exoground_b200.train.step import TanTrainStep  # noqa: E402
from exoground_b200.utils import synth  # noqa: E402

B, T, N, D_IN = 128, 64, 32, 4096
torch.manual_seed(888)
np.random.seed(888)
model = TemporalAligner(6, 6, use_alignability_head=1, video_embed_dim=D_IN, text_embed_dim=D_IN,
                        compute_dtype=torch.bfloat16)
args = synth.tan_args(loss_threshold=0.5, use_alignability_head=1)
tr = TanTrainStep(model, args, B, T, N, D_IN, D_IN, use_graph=False)
batch = synth.tan_batch(B=B, T=T, N=N, d_video=D_IN, d_text=D_IN, zero_pad_text=False)
tr.load_batch(batch)
tr.prepare()
for _ in range(3):
    tr.run()
torch.cuda.synchronize()

K = ops.get()
orig = K.gemm
recs = []


def gemm(A, Bm, **kw):
    a2 = A[0] if A.dim() == 3 else A
    b2 = Bm[0] if Bm.dim() == 3 else Bm
    tA, tB = kw.get("transA", False), kw.get("transB", False)
    M, Kd = (a2.shape[1], a2.shape[0]) if tA else (a2.shape[0], a2.shape[1])
    Nn = b2.shape[1] if tB else b2.shape[0]
    batch = A.shape[0] if A.dim() == 3 else (Bm.shape[0] if Bm.dim() == 3 else 1)
    od = kw.get("out_dtype") or (kw["out"].dtype if kw.get("out") is not None else A.dtype)
    key = (M, Nn, Kd, batch, int(tA), int(tB), kw.get("epilogue", 0), str(od).replace("torch.", ""),
           "acc" if kw.get("out") is not None else "", "pre" if kw.get("want_preact") else "")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    r = orig(A, Bm, **kw)
    e1.record()
    recs.append((key, e0, e1))
    return r


K.gemm = gemm
tr.run()
torch.cuda.synchronize()
K.gemm = orig
agg = collections.OrderedDict()
for key, e0, e1 in recs:
    d = agg.setdefault(key, [0, 0.0])
    d[0] += 1
    d[1] += e0.elapsed_time(e1) * 1e3
tot = sum(v[1] for v in agg.values())
print(f"{'M':>6} {'N':>6} {'K':>6} {'b':>3} tA tB epi {'out':>9} {'':>4} {'n':>3} {'us/call':>8} {'total us':>9} {'TF/s':>7} share")
for key, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    M, Nn, Kd, b, tA, tB, epi, od, acc, pre = key
    fl = 2.0 * M * Nn * Kd * b
    print(f"{M:6d} {Nn:6d} {Kd:6d} {b:3d} {tA:2d} {tB:2d} {epi:3d} {od:>9} {acc + pre:>6} {n:3d} {us / n:8.1f} {us:9.1f} "
          f"{fl / (us / n) / 1e6:7.1f} {100 * us / tot:5.1f}%")
print(f"total {tot:.1f} us over {len(recs)} calls")
