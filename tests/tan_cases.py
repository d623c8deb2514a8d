"""Shared TAN parity case: drop-in TemporalAligner + get_loss + backward against a golden fixture of
the reference (tests/golden/*.npz).  Used by the CPU host-logic test (fake backend, fp32) and by the
GPU parity tests (CUDA kernels, fp32 and bf16)."""
import os

import numpy as np
import torch

from exoground_b200.model import TemporalAligner
from exoground_b200.train import loss as exg_loss
from exoground_b200.utils import synth
from tests.shapes import tan_shapes

SUB = 997


def load_golden(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


def rel_l2(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp(min=1e-30)).item()


def close(a, b, rtol, atol, what=""):
    a = torch.as_tensor(np.asarray(a.detach().cpu() if torch.is_tensor(a) else a), dtype=torch.float32)
    b = torch.as_tensor(np.asarray(b), dtype=torch.float32)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert torch.isfinite(a).all(), f"{what}: non-finite"
    assert torch.allclose(a, b, rtol=rtol, atol=atol), \
        f"{what}: max abs err {(a - b).abs().max().item():.3e} (ref max {b.abs().max().item():.3e})"


def run_tan_case(g, losses_cfg, device, compute_dtype, sub=None, check_grads=True):
    """Tolerances (BASELINE.json north star): fp32 -> logits/loss rel 1e-4, grads rel-L2 1e-3;
    bf16 -> logits/loss 2e-2 (relative to the tensor's scale), grads rel-L2 5e-2."""
    fp32 = compute_dtype == torch.float32
    E, D, B, T, N, d_in, np_seed = [int(x) for x in g["meta"]]
    m = TemporalAligner(E, D, use_alignability_head=1, video_embed_dim=d_in, text_embed_dim=d_in,
                        compute_dtype=compute_dtype)
    m.load_state_dict(synth.fill_state_dict(tan_shapes(E, D, d_in, head=1)))
    m = m.to(device)
    b = synth.tan_batch(B=B, T=T, N=N, d_video=d_in, d_text=d_in)
    video, text = b["video"].to(device), b["text"].to(device)
    vmask, tmask = b["video_padding_mask"].to(device), b["text_padding_mask"].to(device)
    gen = torch.Generator().manual_seed(1234)
    abs_pos = torch.rand(B, N, 2, generator=gen).to(device)
    valid = ~b["text_padding_mask"]
    inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
    for i, (th, al, use_abs) in enumerate(losses_cfg):
        m.zero_grad()
        np.random.seed(np_seed)
        out = m(video, text, vmask, tmask, None)
        args = synth.tan_args(loss_threshold=th, use_alignability_head=al)
        L = exg_loss.get_loss(inp, video, text, vmask, tmask, out, args, abs_pos if use_abs else None)
        for k, v in L.items():
            if not fp32 and (k in ("alignability_top1", "loss-joint-bce") or (k == "loss" and al)):
                continue          # median-split BCE labels / thresholded counts flip under bf16 rounding: not smooth
            close(v, g[f"loss{i}/{k}"], rtol=1e-4 if fp32 else 2e-2, atol=1e-5 if fp32 else 2e-3, what=f"loss{i}/{k}")
        if i == 0 and check_grads:
            # smooth loss (no quantile keep-set, no BCE labels): gradient parity in both precisions
            assert (th, al) == (0.0, 0)
            L["loss"].backward()
            _check_grads(m, g, "grad0", fp32)
            m.zero_grad()
    L["loss"].backward()
    for k, v in out.items():
        a = v.detach().float().cpu()
        ref = torch.as_tensor(g[f"out/{k}"])
        if k.startswith("logits_"):
            if sub:
                a = a[:, :, ::sub[0], :, ::sub[1]]
            vm = valid[:, ::sub[1]] if sub else valid
            a, ref = a[:, :, :, vm], ref[:, :, :, vm]
        elif k == "dual_feature_text":
            a, ref = a[valid], ref[valid]
        elif k == "dual_feature_video" and sub:
            a = a[:, :, ::sub[0], ::16]
        scale = ref.abs().max().item()
        if fp32:
            close(a, ref, rtol=1e-4, atol=1e-4 * scale, what=k)
        else:
            assert (a - ref).abs().max().item() <= 2e-2 * scale + 1e-3, f"{k}: {(a - ref).abs().max().item()} vs scale {scale}"
    if fp32:
        sim = (out["logits_dual"][:, -1] + out["logits_joint"][:, -1]).float().cpu() / 2
        own = torch.stack([sim[bb, :, bb, :] for bb in range(B)], 0)
        assert torch.equal(own.argmax(dim=1)[valid], torch.as_tensor(g["argmax_t"])[valid]), \
            f"argmax-over-time differs (min top-2 gap of the fixture: {float(g['top2_gap_min']):.3e})"
    if check_grads and fp32:
        _check_grads(m, g, "grad", fp32)      # gradients of the last (thresholded / BCE) configuration
    return m


def _check_grads(m, g, tag, fp32, tol_mult=1.0):
    """Per-parameter rel-L2 against the reference's gradients: 1e-3 in fp32, 5e-2 in bf16."""
    for name, p in m.named_parameters():
        if f"{tag}/{name}" not in g:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, name
            continue
        gr = p.grad.detach().float().cpu().reshape(-1)
        gr = gr if gr.numel() <= 4096 else gr[::SUB]
        ref = torch.as_tensor(g[f"{tag}/{name}"])
        e = rel_l2(gr, ref)
        tol = (1e-3 if fp32 else 5e-2) * tol_mult
        if "pos_embed" in name:
            tol *= 4     # a handful of rows with tiny gradients
        assert e <= tol, f"{tag} {name}: rel-L2 {e:.3e} > {tol}"


def run_agree_case(g, device, compute_dtype):
    """Self-labelling branch (train/loss.py:91-232) against tests/golden/tan_agree.npz (oracle/make_golden.py::
    agree_case): the four temporal_agreement_type modes on an 'init' model and one 'cotrain' call."""
    fp32 = compute_dtype == torch.float32
    E, D, B, T, N, d_in, np_seed = [int(x) for x in g["meta"]]
    m = TemporalAligner(E, D, use_alignability_head=1, video_embed_dim=d_in, text_embed_dim=d_in,
                        compute_dtype=compute_dtype)
    m.load_state_dict(synth.fill_state_dict(tan_shapes(E, D, d_in, head=1)))
    m = m.to(device)
    b = synth.tan_batch(B=B, T=T, N=N, d_video=d_in, d_text=d_in, seed=4242, annot_inside_video=True)
    video, text = b["video"].to(device), b["text"].to(device)
    vmask, tmask = b["video_padding_mask"].to(device), b["text_padding_mask"].to(device)
    inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
    tol = dict(rtol=1e-4, atol=1e-5) if fp32 else dict(rtol=2e-2, atol=2e-3)
    for mode in ("keep", "i", "u", "keep-joint"):
        m.zero_grad()
        np.random.seed(np_seed)
        out = m(video, text, vmask, tmask, None)
        args = synth.tan_args(learn_agreement=1, temporal_agreement_type=mode)
        L = exg_loss.get_loss(inp, video, text, vmask, tmask, out, args, None)
        assert {"confidence-ratio", "iou-threshold", "loss", "loss-dual", "loss-joint"} <= set(L)
        if fp32:        # the self-labels are argmax decisions: compared in fp32 only
            for k, v in L.items():
                close(v, g[f"loss_{mode}/{k}"], what=f"{mode}/{k}", **tol)
        else:
            assert torch.isfinite(L["loss"])
        if mode == "keep" and fp32:
            L["loss"].backward()
            # the -6e4 artefacts put terms of 6e4 next to terms of order 1 in the reference's own fp32 sums
            _check_grads(m, g, "grad_keep", True, tol_mult=3.0)
    m.zero_grad()
    np.random.seed(np_seed)
    out = m(video, text, vmask, tmask, None)
    np.random.seed(np_seed + 1)
    with torch.no_grad():
        ema = m(video, text, vmask, tmask, None)
    out["ema-logits_dual"], out["ema-logits_joint"] = ema["logits_dual"], ema["logits_joint"]
    args = synth.tan_args(model="cotrain", learn_agreement=1, temporal_agreement_type="keep")
    L = exg_loss.get_loss(inp, video, text, vmask, tmask, out, args, None)
    if fp32:
        for k, v in L.items():
            close(v, g[f"loss_cotrain/{k}"], what=f"cotrain/{k}", **tol)
