"""Host-side profile of one ExoGround step (cProfile, device synchronised at segment ends)."""
import cProfile
import pstats
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from exoground_b200.model import ExoGroundingTransformer          # noqa: E402
from exoground_b200.train import loss_egoexo4d as exo_loss        # noqa: E402
from exoground_b200.utils import synth                            # noqa: E402

B, T, N = 16, 1024, 32
torch.manual_seed(0); np.random.seed(0)
m = ExoGroundingTransformer(2, 2, video_embed_dim=768, text_embed_dim=4096, use_distill_nce_loss=True,
                            random_pos_start=0, compute_dtype=torch.bfloat16).cuda()
b = synth.exo_batch(B=B, T=T, N=N, d_video=768, d_text=4096, d_ego=768)
b = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in b.items()}
args = synth.exo_args(use_distill_nce_loss=True, same_view_negative=True)
seg = {"fwd": 0.0, "loss": 0.0, "bwd": 0.0}


def step(timing=False):
    for p in m.parameters():
        p.grad = None
    t0 = time.perf_counter()
    out = m(b["video_features"], b["narration_features"], b["video_padding_mask"], b["narration_padding_mask"],
            egocentric_video_embed=b["ego_video_features"][:, 0])
    if timing:
        torch.cuda.synchronize()
    t1 = time.perf_counter()
    L, _ = exo_loss.get_loss(b, out, b["narration_padding_mask"], args)
    if timing:
        torch.cuda.synchronize()
    t2 = time.perf_counter()
    L["loss"].backward()
    if timing:
        torch.cuda.synchronize()
    t3 = time.perf_counter()
    if timing:
        seg["fwd"] += t1 - t0; seg["loss"] += t2 - t1; seg["bwd"] += t3 - t2


for _ in range(3):
    step()
torch.cuda.synchronize()
for _ in range(5):
    step(True)
print({k: round(v / 5 * 1e3, 2) for k, v in seg.items()}, "ms per segment")
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    step()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(35)
