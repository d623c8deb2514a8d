"""Pins the oracle (oracle/*.py) against fixtures produced by the unmodified reference
(oracle/make_golden.py -> tests/golden/*.npz).  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import tan_oracle, loss_oracle, exo_oracle, tfm_oracle
from exoground_b200.utils import synth
from tests.shapes import tan_shapes, exo_shapes, vi_shapes

SUB = 997


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


def _close(a, b, rtol=1e-4, atol=2e-5):
    a = torch.as_tensor(np.asarray(a), dtype=torch.float32)
    b = torch.as_tensor(np.asarray(b), dtype=torch.float32)
    assert a.shape == b.shape, (a.shape, b.shape)
    err = (a - b).abs().max().item() if a.numel() else 0.0
    assert torch.allclose(a, b, rtol=rtol, atol=atol), f"max abs err {err}"


def _tan_run(meta, losses_cfg, g, sub=None):
    E, D, B, T, N, d_in, np_seed = [int(x) for x in meta]
    sd = synth.fill_state_dict(tan_shapes(E, D, d_in, head=1))
    sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    b = synth.tan_batch(B=B, T=T, N=N, d_video=d_in, d_text=d_in)
    gen = torch.Generator().manual_seed(1234)
    abs_pos = torch.rand(B, N, 2, generator=gen)
    np.random.seed(np_seed)
    ps = tan_oracle.draw_pos_starts(T, N)
    out = tan_oracle.tan_forward(sd, E, D, b["video"], b["text"], b["video_padding_mask"],
                                 b["text_padding_mask"], ps, use_alignability_head=1)
    valid = ~b["text_padding_mask"]
    for k, v in out.items():
        a = v.detach()
        if sub and k.startswith("logits_"):
            a = a[:, :, ::sub[0], :, ::sub[1]]
            vm = valid[:, ::sub[1]]
        elif sub and k == "dual_feature_video":
            a = a[:, :, ::sub[0], ::16]
        if k.startswith("logits_"):
            vm = valid[:, ::sub[1]] if sub else valid
            _close(a[:, :, :, vm], g[f"out/{k}"][:, :, :, vm.numpy()])
        elif k == "dual_feature_text":
            _close(a[valid], g[f"out/{k}"][valid.numpy()])
        else:
            _close(a, g[f"out/{k}"], atol=5e-5)
    for i, (th, al, use_abs) in enumerate(losses_cfg):
        L = loss_oracle.tan_loss(out, b["start"], b["end"], b["text_padding_mask"], T, N, th, al,
                                 abs_text_pos=abs_pos if use_abs else None)
        for k, v in L.items():
            _close(float(v), g[f"loss{i}/{k}"], rtol=1e-4, atol=1e-5)
    L["loss"].backward()
    for k, v in sd.items():
        if f"grad/{k}" not in g:
            assert v.grad is None or float(v.grad.abs().max()) == 0.0, k
            continue
        gr = v.grad.reshape(-1)
        _close(gr if gr.numel() <= 4096 else gr[::SUB], g[f"grad/{k}"], rtol=2e-3, atol=2e-6)
    sim = (out["logits_dual"][:, -1] + out["logits_joint"][:, -1]) / 2
    own = torch.stack([sim[bb, :, bb, :] for bb in range(B)], 0)
    am = own.argmax(dim=1)
    assert torch.equal(am[valid], torch.as_tensor(g["argmax_t"])[valid])


def test_circulant_known_answer():
    # train/loss.py:19-20 docstring example — the one known-answer the reference documents
    assert loss_oracle.circulant(torch.tensor([0, 1, 2]), 0).tolist() == [[0, 1, 2], [2, 0, 1], [1, 2, 0]]


def test_tan_small(golden_dir):
    g = _load(golden_dir, "tan_small")
    _tan_run(g["meta"], [(0.0, 0, 0), (0.3, 0, 0), (0.0, 1, 1), (0.5, 1, 0)], g)


def test_tan_cfg1(golden_dir):
    g = _load(golden_dir, "tan_cfg1")
    _tan_run(g["meta"], [(0.0, 0, 0), (0.5, 1, 0)], g, sub=(8, 4))


def test_tan_sim(golden_dir):
    g = _load(golden_dir, "tan_sim")
    sd = synth.fill_state_dict(tan_shapes(2, 2, 32, head=1))
    gen = torch.Generator().manual_seed(77)
    video = torch.randn(2, 24, 32, generator=gen)
    text = torch.randn(2, 5, 32, generator=gen)
    _close(tan_oracle.text_visual_sim_joint(sd, 2, video, text, interpolate_from=16), g["joint_interp16"])
    _close(tan_oracle.text_visual_sim_dual(sd, 2, video, text, interpolate_from=16), g["dual_interp16"])
    _close(tan_oracle.text_visual_sim_joint(sd, 2, video, text), g["joint_p0"])
    _close(tan_oracle.text_visual_sim_dual(sd, 2, video, text), g["dual_p0"])


def test_exo_small(golden_dir):
    g = _load(golden_dir, "exo_small")
    E, D, B, T, N, Dv, Dt, np_seed = [int(x) for x in g["meta"]]
    sd = synth.fill_state_dict(exo_shapes(E, D, Dv, Dt, distill=True))
    sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    b = synth.exo_batch(B=B, T=T, N=N, d_video=Dv, d_text=Dt, d_ego=Dv)
    np.random.seed(np_seed)
    p0, _ = exo_oracle.draw_pos_starts(T)
    out = exo_oracle.exo_forward(sd, E, D, b["video_features"], b["narration_features"],
                                 b["video_padding_mask"], b["narration_padding_mask"], p0,
                                 use_distill=True, ego=b["ego_video_features"][:, 0])
    for k in ("interval_preds", "low_dim_features", "high_dim_features"):
        _close(out[k].detach(), g[f"out/{k}"], atol=5e-5)
    G, iou = exo_oracle.grounding_loss(out["interval_preds"], b["mean"], b["duration"], b["narration_padding_mask"])
    xl = exo_oracle.xview_info_nce(out["high_dim_features"], b["ego_video_features"], b["view_rank_label"],
                                   b["view_rank_neg_label"], b["same_view_neg_idxs"], b["video_features"])
    l1, pc, nc = exo_oracle.l1_cosine(out["high_dim_features"], b["ego_video_features"],
                                      b["view_rank_label"], b["view_rank_neg_label"])
    _close(iou.detach(), g["iou"])
    for k in ("Center L1 loss", "Duration L1 loss", "IoU loss", "mean IoU", "IoU>=0.1"):
        _close(float(G[k]), g[f"loss/{k}"])
    _close(float(xl.mean()), g["loss/InfoNCE loss"])
    _close(float(l1.mean()), g["loss/L1 loss"])
    _close(float(pc.mean()), g["loss/Pos cosine sim"])
    _close(float(nc.mean()), g["loss/Avg neg cosine sim"])
    loss = G["loss"] + xl.mean()
    _close(float(loss), g["loss/loss"])
    loss.backward()
    for k, v in sd.items():
        if f"grad/{k}" not in g:
            continue
        gr = v.grad.reshape(-1)
        _close(gr if gr.numel() <= 4096 else gr[::SUB], g[f"grad/{k}"], rtol=2e-3, atol=2e-6)
    G2, _ = exo_oracle.grounding_loss(out["interval_preds"].detach(), b["starts"], b["ends"],
                                      b["narration_padding_mask"], use_center_duration=False)
    xl2 = exo_oracle.xview_info_nce(out["high_dim_features"].detach(), b["ego_video_features"],
                                    b["view_rank_label"], b["view_rank_neg_label"])
    _close(float(G2["loss"] + xl2.mean()), g["loss_se/loss"])


def test_keystep_small(golden_dir):
    g = _load(golden_dir, "keystep_small")
    gx = _load(golden_dir, "exo_small")
    E, D, B, T, N, Dv, Dt, np_seed = [int(x) for x in gx["meta"]]
    visd = synth.fill_state_dict(vi_shapes(Dv), seed=7)
    sd = synth.fill_state_dict(exo_shapes(E, D, Dv, Dt, distill=True))
    b = synth.exo_batch(B=B, T=T, N=N, d_video=Dv, d_text=Dt, d_ego=Dv)
    _close(exo_oracle.vi_mlp_forward(visd, b["video_features"]), g["vi/high_dim_features"], atol=5e-5)
    np.random.seed(np_seed)
    p0, _ = exo_oracle.draw_pos_starts(T)
    out = exo_oracle.exo_forward(sd, E, D, b["video_features"], b["narration_features"],
                                 b["video_padding_mask"], b["narration_padding_mask"], p0,
                                 use_distill=True, ego=b["ego_video_features"][:, 0], vi_sd=visd)
    _close(out["interval_preds"], g["out/interval_preds"], atol=5e-5)
    _close(float(out["distill_infonce_loss"]), g["out/distill_infonce_loss"])


def test_interp_matches_torch():
    import torch.nn.functional as F
    tab = torch.randn(40, 8)
    for k, size in [(16, 24), (40, 7), (10, 10), (5, 64)]:
        ref = F.interpolate(tab[None, :k].transpose(1, 2), size=size, mode="linear",
                            align_corners=False).transpose(1, 2)[0]
        _close(tfm_oracle.interp_pos(tab, k, size), ref, atol=1e-6)
