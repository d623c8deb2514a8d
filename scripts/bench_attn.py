"""Attention micro-benchmark through the C ABI: TFLOP/s of exg_attn_fwd / exg_attn_bwd (bf16, 8 heads x 64) over the
sequence lengths of the T = 64 -> 4096 inference sweep (BASELINE.json configs[4]) and ExoGround's T = 1024."""
import sys

import torch

sys.path.insert(0, ".")
from exoground_b200 import ops  # noqa: E402

K = ops.get()
H, d = 8, 64
print(f"{'B':>4} {'S':>6} {'fwd ms':>9} {'fwd TF/s':>9} {'bwd ms':>9} {'bwd TF/s':>9}")
only = [int(x) for x in sys.argv[1:]]
# (S, B): the train step's own shapes first (video encoder T=64, joint encoder T+N=96, 128 clips), then the sweep
for S, B in [(64, 128), (96, 128), (32, 128)] + [(S, max(1, 16384 // S)) for S in (64, 128, 256, 512, 1024, 2048, 4096)]:
    if only and S not in only:
        continue
    qkv = torch.randn(B * S, 3 * H * d, device="cuda").bfloat16()
    q, k, v = qkv[:, :512], qkv[:, 512:1024], qkv[:, 1024:]
    scale = d ** -0.5
    o, lse = K.attn_fwd(q, k, v, None, B, H, S, S, scale)
    do = torch.randn_like(o)
    dqkv = torch.zeros_like(qkv)

    def fwd():
        K.attn_fwd(q, k, v, None, B, H, S, S, scale)

    def bwd():
        K.attn_bwd(q, k, v, o, do, lse, None, dqkv[:, :512], dqkv[:, 512:1024], dqkv[:, 1024:], B, H, S, S, scale)

    res = []
    for fn, fl in ((fwd, 4.0), (bwd, 10.0)):
        # CUDA-graph replay of 10 launches: the eager launch path (ctypes + 4-7 tensor-map encodes per call, ~25 us of
        # host time) would hide every kernel shorter than that
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        n = 10
        with torch.cuda.graph(g):
            for _ in range(n):
                fn()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / (n * reps)
        res += [ms, fl * B * H * S * S * d / (ms * 1e-3) / 1e12]
    print(f"{B:>4} {S:>6} {res[0]:>9.4f} {res[1]:>9.1f} {res[2]:>9.4f} {res[3]:>9.1f}", flush=True)
