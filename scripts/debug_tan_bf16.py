"""GPU debug helper: bf16 TAN parity numbers vs the golden fixture, printed instead of asserted."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from exoground_b200.model import TemporalAligner
from exoground_b200.train import loss as exg_loss
from exoground_b200.utils import synth
from tests.shapes import tan_shapes
from tests.tan_cases import load_golden, rel_l2, SUB

name = sys.argv[1] if len(sys.argv) > 1 else "tan_cfg1"
dt = torch.float32 if (len(sys.argv) > 2 and sys.argv[2] == "fp32") else torch.bfloat16
g = load_golden(os.path.join(os.path.dirname(__file__), "..", "tests", "golden"), name)
E, D, B, T, N, d_in, np_seed = [int(x) for x in g["meta"]]
m = TemporalAligner(E, D, use_alignability_head=1, video_embed_dim=d_in, text_embed_dim=d_in, compute_dtype=dt)
m.load_state_dict(synth.fill_state_dict(tan_shapes(E, D, d_in, head=1)))
m = m.to("cuda")
b = synth.tan_batch(B=B, T=T, N=N, d_video=d_in, d_text=d_in)
dev = lambda t: t.to("cuda")
video, text, vmask, tmask = dev(b["video"]), dev(b["text"]), dev(b["video_padding_mask"]), dev(b["text_padding_mask"])
abs_pos = torch.rand(B, N, 2, generator=torch.Generator().manual_seed(1234)).to("cuda")
inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
cfgs = [(0.0, 0, 0), (0.3, 0, 0), (0.0, 1, 1), (0.5, 1, 0)] if name == "tan_small" else [(0.0, 0, 0), (0.5, 1, 0)]
for i, (th, al, ua) in enumerate(cfgs):
    m.zero_grad()
    np.random.seed(np_seed)
    out = m(video, text, vmask, tmask, None)
    L = exg_loss.get_loss(inp, video, text, vmask, tmask, out, synth.tan_args(loss_threshold=th, use_alignability_head=al),
                          abs_pos if ua else None)
    for k, v in L.items():
        ref = float(g[f"loss{i}/{k}"])
        print(f"loss{i}/{k}: got {float(v):.6f} ref {ref:.6f} rel {abs(float(v)-ref)/max(abs(ref),1e-9):.3e}")
    if i == 0:
        break
L["loss"].backward()
errs = []
for n, p in m.named_parameters():
    if f"grad0/{n}" not in g or p.grad is None:
        continue
    gr = p.grad.detach().float().cpu().reshape(-1)
    gr = gr if gr.numel() <= 4096 else gr[::SUB]
    errs.append((rel_l2(gr, torch.as_tensor(g[f"grad0/{n}"])), n))
for e, n in sorted(errs, reverse=True)[:25]:
    print(f"grad rel-L2 {e:.3e}  {n}")
print("median grad rel-L2", sorted(errs)[len(errs)//2][0])
