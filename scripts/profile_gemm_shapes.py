"""Per-shape GEMM time inside one EAGER train step (CUDA events around every exg_gemm call), for the TAN headline step
(default) or the ExoGround E6D6 T=1024 step (--exo): which shapes of the step run far below the class average.

    python scripts/profile_gemm_shapes.py [--exo] [--steps 3]
"""
import argparse
import collections
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from exoground_b200 import ops
from exoground_b200.utils import synth


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--exo", action="store_true")
    ap.add_argument("--steps", type=int, default=3)
    a = ap.parse_args()
    K = ops.get()
    torch.manual_seed(0); np.random.seed(0)
    if a.exo:
        from exoground_b200.model import ExoGroundingTransformer
        from exoground_b200.train.step_exo import ExoTrainStep
        B, T, N = 16, 1024, 64
        m = ExoGroundingTransformer(6, 6, video_embed_dim=768, text_embed_dim=4096, random_pos_start=0,
                                    compute_dtype=torch.bfloat16)
        args = synth.exo_args()
        st = ExoTrainStep(m, args, use_graph=False)
        batch = synth.exo_batch(B=B, T=T, N=N, d_video=768, d_text=4096, d_ego=768)
        step = lambda: st.step(batch)
    else:
        from exoground_b200.model import TemporalAligner
        from exoground_b200.train.step import TanTrainStep
        B, T, N, D = 128, 64, 32, 4096
        m = TemporalAligner(6, 6, use_alignability_head=1, video_embed_dim=D, text_embed_dim=D,
                            compute_dtype=torch.bfloat16)
        args = synth.tan_args(loss_threshold=0.5, use_alignability_head=1)
        st = TanTrainStep(m, args, B, T, N, D, D, use_graph=False, branch_streams=False)
        batch = synth.tan_batch(B=B, T=T, N=N, d_video=D, d_text=D, zero_pad_text=False)
        st.load_batch(batch)
        st.prepare()
        step = lambda: st.run()
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    rec = []
    orig = K.gemm

    def timed(A, Bm, *args_, **kw):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = orig(A, Bm, *args_, **kw)
        e1.record()
        tA, tB = bool(kw.get("transA", False)), bool(kw.get("transB", False))
        a2, b2 = (A[0] if A.dim() == 3 else A), (Bm[0] if Bm.dim() == 3 else Bm)
        nb = A.shape[0] if A.dim() == 3 else (Bm.shape[0] if Bm.dim() == 3 else 1)
        Mm, Kk = (a2.shape[1], a2.shape[0]) if tA else (a2.shape[0], a2.shape[1])
        Nn = b2.shape[1] if tB else b2.shape[0]
        Mm *= nb                                   # batched: report the stacked problem
        key = (Mm, Nn, Kk, int(tA), int(tB), int(kw.get("epilogue", 0) or 0), int(kw.get("bias") is not None),
               str(A.dtype).split(".")[-1], str(kw.get("out_dtype") or "").split(".")[-1])
        rec.append((key, e0, e1))
        return out
    K.gemm = timed
    for _ in range(a.steps):
        step()
    torch.cuda.synchronize()
    K.gemm = orig
    agg = collections.OrderedDict()
    for key, e0, e1 in rec:
        d = agg.setdefault(key, [0, 0.0])
        d[0] += 1
        d[1] += e0.elapsed_time(e1)
    tot = sum(v[1] for v in agg.values()) / a.steps
    print(f"# {'ExoGround E6D6 T=1024 N=64 B=16' if a.exo else 'TAN E6D6 T=64 N=32 B=128'}: {len(rec) // a.steps} GEMM calls / step, "
          f"{tot:.3f} ms / step in GEMMs (eager, events around each call incl. launch gaps)")
    print(f"{'M':>6} {'N':>6} {'K':>6} tA tB epi bias {'in':>8} {'out':>8} {'calls':>5} {'us/call':>8} {'TF/s':>7} {'ms/step':>8} {'share':>6}")
    for key, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        Mm, Nn, Kk = key[:3]
        us = ms / n * 1e3
        print(f"{Mm:>6} {Nn:>6} {Kk:>6} {key[3]:>2} {key[4]:>2} {key[5]:>3} {key[6]:>4} {key[7]:>8} {key[8]:>8} {n // a.steps:>5} {us:>8.1f} "
              f"{2.0 * Mm * Nn * Kk / us / 1e6:>7.0f} {ms / a.steps:>8.3f} {ms / a.steps / tot:>6.1%}")


if __name__ == "__main__":
    main()
