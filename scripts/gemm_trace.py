"""GPU helper: pipeline timeline of one CTA of the persistent tcgen05 GEMM (exg_gemm_trace) on the shapes of the TAN
train step.  Prints, per shape: kernel time (CUDA events), and for the traced CTA the cycle stamps of setup, first
operands, every tile's MMA span and every epilogue warp's chunk phases, so one can see which role the tile time
belongs to (TMA feed, MMA issue, TMEM drain, staging-buffer wait, store issue).

    python scripts/gemm_trace.py [--cta 0] [--only qkv,c_fc]
"""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from exoground_b200 import ops
from exoground_b200.ops import lib as L

TAGS = {1: "slot_free", 2: "acc_free", 3: "ops_landed", 4: "tile_commit", 5: "acc_ready", 6: "tmem_read", 7: "math",
        8: "buf_free", 9: "store_issued", 10: "tile_done", 11: "entry", 12: "setup_done", 13: "roles_done",
        14: "teardown", 15: "stores_drained"}
SHAPES = [  # (M, N, K, tA, tB, epilogue, bias, out_f32, tag, want_preact)
    (8192, 1536, 512, 0, 0, L.EPI_NONE, 1, 0, "qkv", 0),
    (8192, 512, 512, 0, 0, L.EPI_NONE, 1, 0, "out_proj", 0),
    (12288, 2048, 512, 0, 0, L.EPI_QGELU, 1, 0, "c_fc+qgelu+pre", 1),
    (12288, 512, 2048, 0, 0, L.EPI_NONE, 1, 0, "c_proj", 0),
    (12288, 2048, 512, 0, 1, L.EPI_NONE, 0, 0, "du", 0),
    (12288, 2048, 512, 0, 1, L.EPI_QGELU_BWD, 0, 0, "du+qgelu_bwd", 0),
    (2048, 512, 12288, 1, 1, L.EPI_NONE, 0, 1, "dW_fc", 0),
    (512, 512, 8192, 1, 1, L.EPI_NONE, 0, 1, "dW_out", 0),
    (8192, 4096, 512, 0, 0, L.EPI_NONE, 0, 0, "sim", 0),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cta", type=int, default=0)
    ap.add_argument("--only", default="")
    ap.add_argument("--cap", type=int, default=512)
    a = ap.parse_args()
    K = ops.get()
    lib = L.load()
    shapes = [s for s in SHAPES if not a.only or any(k in s[8] for k in a.only.split(","))]
    buf = torch.zeros(11, a.cap, dtype=torch.int64, device="cuda")
    for (M, N, Kd, tA, tB, epi, bias, f32, tag, pre) in shapes:
        A = torch.randn((Kd, M) if tA else (M, Kd), device="cuda").bfloat16()
        B = torch.randn((Kd, N) if tB else (N, Kd), device="cuda").bfloat16()
        bs = torch.randn(N, device="cuda") if bias else None
        out = torch.empty(M, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
        aux = torch.randn(M, N, device="cuda").bfloat16() if epi == L.EPI_QGELU_BWD else None
        kw = dict(transA=bool(tA), transB=bool(tB), bias=bs, epilogue=epi, out=out, aux=aux, want_preact=bool(pre))
        for _ in range(3):
            K.gemm(A, B, **kw)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            K.gemm(A, B, **kw)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 20 * 1e3
        buf.zero_()
        lib.exg_gemm_trace(C.c_void_p(buf.data_ptr()), a.cap, a.cta)
        K.gemm(A, B, **kw)
        torch.cuda.synchronize()
        lib.exg_gemm_trace(None, 0, 0)
        ev = buf.cpu().numpy().astype("uint64")
        rows = {}
        for r in range(11):
            rows[r] = [(int(x) >> 8, int(x) & 255) for x in ev[r] if x != 0]
        t0 = rows[10][0][0] if rows[10] else 0
        print(f"=== {tag}  M{M} N{N} K{Kd}: {us:.1f} us  ({2 * M * N * Kd / us / 1e6:.0f} TF/s)   [traced CTA {a.cta}, cycles from kernel entry]")
        cta = {TAGS[t]: c - t0 for c, t in rows[10]}
        print("  cta     :", " ".join(f"{k}={v}" for k, v in cta.items()))
        prod = [c - t0 for c, t in rows[0]]
        if prod:
            print(f"  producer: {len(prod)} k-blocks issued; first {prod[:10]} ... last {prod[-3:]}")
        tiles, cur = [], None
        for c, t in rows[1]:
            if t == 2:
                cur = {"acc_free": c - t0, "landed": []}
            elif t == 3 and cur is not None:
                cur["landed"].append(c - t0)
            elif t == 4 and cur is not None:
                cur["commit"] = c - t0
                tiles.append(cur)
        for i, tl in enumerate(tiles[:8]):
            ld = tl["landed"]
            print(f"  mma tile {i}: acc_free={tl['acc_free']} first_ops={ld[0]} last_ops={ld[-1]} commit_issued={tl['commit']}"
                  f"  (k-block gaps: {[ld[j + 1] - ld[j] for j in range(min(len(ld) - 1, 9))]})")
        if len(tiles) > 8:
            print(f"  ... {len(tiles)} tiles; tile starts: {[tl['acc_free'] for tl in tiles]}")
        for w in (2, 6):
            seq = [(TAGS[t], c - t0) for c, t in rows[w]]
            # split per tile
            out_lines, line = [], []
            for name, c in seq:
                line.append(f"{name}={c}")
                if name in ("tile_done", "stores_drained"):
                    out_lines.append(" ".join(line)); line = []
            for i, l in enumerate(out_lines[:6]):
                print(f"  epi warp {w - 2} tile {i}: {l}")
            if len(out_lines) > 6:
                print(f"  epi warp {w - 2}: ... {len(out_lines)} records, last: {out_lines[-1]}")
        sys.stdout.flush()


if __name__ == "__main__":
    main()
