#!/usr/bin/env python
"""Does the gradient all-reduce overlap backward?  torchrun --nproc-per-node N scripts/trace_overlap.py [--no-overlap]

Records a kineto trace of a few replayed TAN steps on rank 0 and reports, per step: wall time of the step, total NCCL
kernel time, the part of it during which one of OUR kernels was running on another stream (overlapped) and the part
during which nothing else ran (exposed)."""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from exoground_b200.model import TemporalAligner  # noqa: E402
from exoground_b200.train.step import TanTrainStep  # noqa: E402
from exoground_b200.utils import synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--no-overlap", action="store_true")   # default here: overlapped buckets
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--bucket-mb", type=float, default=25.0)
    a = ap.parse_args()
    rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    B, T, N, D = 128, 64, 32, 4096
    torch.manual_seed(888)
    np.random.seed(888 + rank)
    model = TemporalAligner(6, 6, use_alignability_head=1, video_embed_dim=D, text_embed_dim=D,
                            compute_dtype=torch.bfloat16)
    args = synth.tan_args(loss_threshold=0.5, use_alignability_head=1)
    tr = TanTrainStep(model, args, B, T, N, D, D, device=f"cuda:{local}", process_group=dist.group.WORLD,
                      overlap_allreduce=not a.no_overlap, bucket_mb=a.bucket_mb)
    tr.load_batch(synth.tan_batch(B=B, T=T, N=N, d_video=D, d_text=D, seed=888 + rank, zero_pad_text=False))
    tr.prepare()
    for _ in range(5):
        tr.run()
    torch.cuda.synchronize()
    dist.barrier()
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(a.steps):
            tr.run()
        torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range.end > e.time_range.start]
        ev = [(e.time_range.start, e.time_range.end, e.name) for e in ev if "Memcpy" not in e.name and "Memset" not in e.name]
        ev.sort()
        nccl = [(s, t) for s, t, n in ev if "nccl" in n.lower()]
        ours = sorted((s, t) for s, t, n in ev if "nccl" not in n.lower())
        # union of our kernels' intervals
        uni = []
        for s, t in ours:
            if uni and s <= uni[-1][1]:
                uni[-1][1] = max(uni[-1][1], t)
            else:
                uni.append([s, t])
        def covered(s, t):
            c = 0.0
            for us, ut in uni:
                if ut <= s:
                    continue
                if us >= t:
                    break
                c += min(t, ut) - max(s, us)
            return c
        tot = sum(t - s for s, t in nccl)
        ov = sum(covered(s, t) for s, t in nccl)
        span = (ev[-1][1] - ev[0][0]) / a.steps
        busy = sum(t - s for s, t in uni) / a.steps
        print(f"mode={'serial' if a.no_overlap else 'overlapped'} steps={a.steps} nccl_kernels/step={len(nccl) / a.steps:.1f} "
              f"step_span={span / 1e3:.3f} ms  own_kernels_busy={busy / 1e3:.3f} ms  nccl_total={tot / a.steps / 1e3:.3f} ms  "
              f"nccl_overlapped_with_compute={ov / a.steps / 1e3:.3f} ms  nccl_exposed={(tot - ov) / a.steps / 1e3:.3f} ms")
        for s, t in nccl[:len(nccl) // a.steps]:
            print(f"  nccl kernel: start +{(s - ev[0][0]) / 1e3:.3f} ms  dur {(t - s) / 1e3:.3f} ms  overlapped {covered(s, t) / 1e3:.3f} ms")
    tr.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
