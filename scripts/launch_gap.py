"""Per-node cost of dependent tiny kernels inside a CUDA graph (launch gap measurement)."""
import torch
x = torch.zeros(4, device="cuda")
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    for _ in range(3):
        x.add_(1)
torch.cuda.synchronize()
for n in (100, 400):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n):
            x.add_(1)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    print(n, "nodes:", e0.elapsed_time(e1) / 10 / n * 1e3, "us per node")
