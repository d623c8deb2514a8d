"""Runs the same 4 training steps several times (fresh model each time) and prints the losses."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from exoground_b200.model import TemporalAligner
from exoground_b200.train.step import TanTrainStep
from exoground_b200.utils import synth
from tests.shapes import tan_shapes

from tests.test_train_step_gpu import B, T, N, DIN, E, D
args = synth.tan_args(loss_threshold=0.0, use_alignability_head=0)
batches = [synth.tan_batch(B=B, T=T, N=N, d_video=DIN, d_text=DIN, seed=300 + i) for i in range(4)]
for rep in range(4):
    m = TemporalAligner(E, D, use_alignability_head=1, video_embed_dim=DIN, text_embed_dim=DIN, compute_dtype=torch.bfloat16)
    m.load_state_dict(synth.fill_state_dict({k: v.shape for k, v in m.state_dict().items()}))
    m.random_pos_start = 0
    tr = TanTrainStep(m, args, B, T, N, DIN, DIN, lr=1e-3, eps=float(os.environ.get('EPS', '1e-8')), use_graph=(rep % 2 == 0))
    tr.load_batch(batches[0]); tr.prepare()
    print(rep, [tr.step(b) for b in batches])
