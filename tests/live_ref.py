"""Shared bodies of the LIVE-reference parity tests: the drop-in (CUDA kernels through the C ABI, or the op-spec
backend on CPU) against the UNMODIFIED reference modules run in the same process on the same device.

The reference comes from /root/reference (build container) or from the packed archive oracle/_ref (GPU box);
see oracle/ref_import.py.  Tolerances are BASELINE.json's: fp32 logits / loss rel 1e-4, gradients rel-L2 1e-3,
argmax over time bit-exact; bf16 logits / loss 2e-2, gradients rel-L2 5e-2.

bf16 and the DISCRETE parts of the TAN loss (quantile keep-set, median-split alignability labels, top-1): those are
step functions of the logits, so a bf16 logit that differs from the fp32 one in the third digit can legitimately flip
a label.  Instead of skipping those keys the bf16 tests split the check in two, each of which is tight:
  (1) logits of the drop-in (bf16 compute) vs the reference's fp32 logits: 2e-2 of the tensor's scale;
  (2) the reference's OWN get_loss evaluated on the drop-in's logits vs the drop-in's get_loss on the same logits:
      every key (BCE, top-1, thresholded means included) to 1e-3 - the labels are then forced to be the same.
"""
import numpy as np
import torch

from exoground_b200.model import TemporalAligner, ExoGroundingTransformer
from exoground_b200.train import loss as exg_loss
from exoground_b200.train import loss_egoexo4d as exo_loss
from exoground_b200.utils import synth
from oracle import ref_import


def rel_l2(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp(min=1e-30)).item()


def tan_pair(E, D, d_in, head, device, dtype, seed=888, **kw):
    """(reference module namespace, reference TemporalAligner fp32, drop-in TemporalAligner) with equal weights."""
    r = ref_import.load()
    with ref_import.quiet():
        ref = r.tan_model.TemporalAligner(E, D, use_alignability_head=head, **kw)
    if d_in != 4096:        # hard-coded 4096 in the reference (tan_model.py:42-43)
        ref.video_pre_proj = torch.nn.Linear(d_in, 512, bias=False)
        ref.text_pre_proj = torch.nn.Linear(d_in, 512, bias=False)
    sd = synth.fill_state_dict({k: v.shape for k, v in ref.state_dict().items()}, seed=seed)
    ref.load_state_dict(sd)
    own = TemporalAligner(E, D, use_alignability_head=head, video_embed_dim=d_in, text_embed_dim=d_in,
                          compute_dtype=dtype, **kw)
    own.load_state_dict(sd)
    return r, ref.to(device), own.to(device)


def _to(b, device):
    return {k: (v.to(device) if torch.is_tensor(v) else v) for k, v in b.items()}


def _float_dict(L):
    return {k: float(v) for k, v in L.items() if not isinstance(v, dict)}


def check_grads(own, ref, fp32, tol_mult=1.0, skip=("in_proj_bias",), what=""):
    """Per-parameter rel-L2 of the drop-in's gradients against the reference's.  `in_proj_bias`: its key third has
    an identically-zero gradient (softmax shift invariance) that is rounding noise on both sides - compare q, v."""
    gr = {n: p.grad for n, p in ref.named_parameters() if p.grad is not None}
    go = {n: p.grad for n, p in own.named_parameters() if p.grad is not None}
    missing = [n for n in gr if n not in go and float(gr[n].abs().max()) > 0]
    assert not missing, f"{what}: parameters without a gradient in the drop-in: {missing[:5]}"
    tol = (1e-3 if fp32 else 5e-2) * tol_mult
    worst = (0.0, "")
    for n, g in gr.items():
        if n not in go or float(g.norm()) < 1e-9:
            continue
        a, c = go[n].detach().float().reshape(-1), g.detach().float().reshape(-1)
        if n.endswith(skip):
            keep = torch.cat((torch.arange(0, 512), torch.arange(1024, 1536))).to(a.device)
            a, c = a[keep], c[keep]
        e = rel_l2(a, c)
        t = tol * (4 if "pos_embed" in n else 1)
        assert e <= t, f"{what} grad {n}: rel-L2 {e:.3e} > {t:.1e}"
        worst = max(worst, (e, n))
    return worst


def run_tan_live(device, dtype, E, D, B, T, N, d_in, th, head, *, seed=888, np_seed=5, batch_kw=None,
                 push_annotation=False, grads=True, model_kw=None, what="tan"):
    """Drop-in TemporalAligner + get_loss + backward against the live reference on `device`."""
    fp32 = dtype == torch.float32
    r, ref, own = tan_pair(E, D, d_in, head, device, dtype, seed=seed, **(model_kw or {}))
    kw = dict(zero_pad_text=False)
    kw.update(batch_kw or {})
    b = synth.tan_batch(B=B, T=T, N=N, d_video=d_in, d_text=d_in, seed=seed, **kw)
    if push_annotation:      # an annotation past the end of the clip: rows and a column without any positive
        b["start"][-1][-1], b["end"][-1][-1] = float(T + 3), float(T + 5)
    bd = _to(b, device)
    inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
    valid = ~bd["text_padding_mask"]
    args = synth.tan_args(loss_threshold=th, use_alignability_head=head)
    smooth = synth.tan_args(loss_threshold=0.0, use_alignability_head=0)
    report = {}

    def fwd(mod):
        np.random.seed(np_seed)
        return mod(bd["video"], bd["text"], bd["video_padding_mask"], bd["text_padding_mask"], None)

    with ref_import.quiet():
        o_r = fwd(ref)
        L_r = r.loss.get_loss(inp, bd["video"], bd["text"], bd["video_padding_mask"], bd["text_padding_mask"], o_r,
                              args, None)
    o_o = fwd(own)
    L_o = exg_loss.get_loss(inp, bd["video"], bd["text"], bd["video_padding_mask"], bd["text_padding_mask"], o_o,
                            args, None)
    # ---- outputs
    assert set(o_o) == set(o_r), (set(o_o), set(o_r))
    for k in o_r:
        a, c = o_o[k].detach().float(), o_r[k].detach().float()
        assert a.shape == c.shape, (k, a.shape, c.shape)
        if k.startswith("logits_"):
            a, c = a[:, :, :, valid], c[:, :, :, valid]          # padded text columns: reference trap (0/0 rows)
        elif k == "dual_feature_text":
            a, c = a[valid], c[valid]
        elif "alignability" in k:
            a, c = a[valid.unsqueeze(1).expand(a.shape[:-1])] if a.dim() == 4 else a[valid], \
                   c[valid.unsqueeze(1).expand(c.shape[:-1])] if c.dim() == 4 else c[valid]
        scale = float(c.abs().max())
        err = float((a - c).abs().max())
        tol = 1e-4 if fp32 else 2e-2
        assert err <= tol * scale + (1e-6 if fp32 else 1e-3), f"{what} {k}: max abs err {err:.3e}, scale {scale:.3e}"
        report[k] = err / max(scale, 1e-30)
    # ---- loss dict
    assert set(L_o) == set(L_r), (set(L_o), set(L_r))
    fr, fo = _float_dict(L_r), _float_dict(L_o)
    if fp32:
        for k in fr:
            assert abs(fo[k] - fr[k]) <= 1e-4 * max(1.0, abs(fr[k])), (what, k, fo[k], fr[k])
    else:
        # (2) of the module docstring: same logits -> the reference's own loss code against ours, every key
        with ref_import.quiet(), torch.no_grad():
            same = {k: (v.detach().float() if torch.is_tensor(v) else v) for k, v in o_o.items()}
            L_same = _float_dict(r.loss.get_loss(inp, bd["video"], bd["text"], bd["video_padding_mask"],
                                                 bd["text_padding_mask"], same, args, None))
        for k in L_same:
            assert abs(fo[k] - L_same[k]) <= 1e-3 * max(1.0, abs(L_same[k])), (what, "same-logits", k, fo[k], L_same[k])
        # and the smooth keys directly against the fp32 reference
        for k in ("loss-dual-all", "loss-joint-all", "loss-total") if th > 0 else ("loss-dual", "loss-joint"):
            if k in fr:
                assert abs(fo[k] - fr[k]) <= 2e-2 * max(1.0, abs(fr[k])), (what, k, fo[k], fr[k])
    report["loss"] = (fo["loss"], fr["loss"])
    # ---- predicted timestamps (fp32: bit-exact), with the smallest top-2 gap of this synthetic set
    if fp32:
        s_o = (o_o["logits_dual"][:, -1] + o_o["logits_joint"][:, -1]).float() / 2
        s_r = (o_r["logits_dual"][:, -1] + o_r["logits_joint"][:, -1]).float() / 2
        idx = torch.arange(B, device=s_o.device)
        own_o, own_r = s_o[idx, :, idx], s_r[idx, :, idx]            # [B,T,N]
        assert torch.equal(own_o.argmax(1)[valid], own_r.argmax(1)[valid]), f"{what}: argmax over time differs"
        top2 = own_r.topk(2, dim=1).values
        report["top2_gap_min"] = float((top2[:, 0] - top2[:, 1])[valid].min())
    # ---- gradients
    if grads:
        if fp32:
            L_r["loss"].backward()
            L_o["loss"].backward()
            report["grad_worst"] = check_grads(own, ref, True, what=what)
        else:
            # smooth configuration (no keep-set / labels): the bf16 gradient against the fp32 reference
            with ref_import.quiet():
                Ls_r = r.loss.get_loss(inp, bd["video"], bd["text"], bd["video_padding_mask"], bd["text_padding_mask"],
                                       o_r, smooth, None)
            Ls_o = exg_loss.get_loss(inp, bd["video"], bd["text"], bd["video_padding_mask"], bd["text_padding_mask"],
                                     o_o, smooth, None)
            Ls_r["loss"].backward()
            Ls_o["loss"].backward()
            report["grad_worst"] = check_grads(own, ref, False, what=what)
    return report


def exo_pair(E, D, Dv, Dt, device, dtype, seed=41, **kw):
    r = ref_import.load()
    with ref_import.quiet():
        ref = r.exo_ground_model.ExoGroundingTransformer(E, D, video_embed_dim=Dv, text_embed_dim=Dt, **kw)
    sd = synth.fill_state_dict({k: v.shape for k, v in ref.state_dict().items()}, seed=seed)
    ref.load_state_dict(sd)
    own = ExoGroundingTransformer(E, D, video_embed_dim=Dv, text_embed_dim=Dt, compute_dtype=dtype, **kw)
    own.load_state_dict(sd)
    return r, ref.to(device), own.to(device)


def run_exo_live(device, dtype, E, D, B, T, N, Dv, Dt, *, distill, cd=True, svn=False, only=False, dec=True, tpos=0,
                 seed=1, min_text=4, np_seed=5, what="exo", grad_tol_mult=1.0, random_pos_start=1):
    """ExoGroundingTransformer + loss_egoexo4d.get_loss + backward against the live reference on `device`."""
    fp32 = dtype == torch.float32
    r, ref, own = exo_pair(E, D, Dv, Dt, device, dtype, use_decoder=dec, use_distill_nce_loss=distill,
                           use_text_pos_enc=tpos, random_pos_start=random_pos_start)
    b = synth.exo_batch(B=B, T=T, N=N, d_video=Dv, d_text=Dt, d_ego=Dv if distill else None, seed=seed,
                        min_text=min_text)
    bd = _to(b, device)
    args = synth.exo_args(use_distill_nce_loss=distill, use_center_duration=cd, same_view_negative=svn,
                          only_same_view_negative=only)
    outs = []
    for mod, get_loss in ((ref, r.loss_egoexo4d.get_loss), (own, exo_loss.get_loss)):
        np.random.seed(np_seed)
        with ref_import.quiet():
            out = mod(bd["video_features"], bd["narration_features"], bd["video_padding_mask"],
                      bd["narration_padding_mask"],
                      egocentric_video_embed=bd["ego_video_features"][:, 0] if distill else None)
            if not dec:
                out = dict(out, interval_preds=out["interval_preds"][:, -1])
            L, iou = get_loss(bd, out, bd["narration_padding_mask"], args)
        L["loss"].backward()
        outs.append((out, L, iou))
    (o_r, L_r, i_r), (o_o, L_o, i_o) = outs
    report = {}
    for k in o_r:
        a, c = o_o[k].detach().float(), o_r[k].detach().float()
        scale = float(c.abs().max())
        err = float((a - c).abs().max())
        assert a.shape == c.shape and err <= (1e-4 if fp32 else 2e-2) * scale + (1e-6 if fp32 else 1e-3), \
            f"{what} {k}: max abs err {err:.3e}, scale {scale:.3e}"
        report[k] = err / max(scale, 1e-30)
    if fp32:
        assert torch.allclose(i_o.detach().float().cpu(), i_r.detach().float().cpu(), rtol=1e-4, atol=1e-5)
    for k, v in L_r.items():
        if isinstance(v, dict):
            a, c = float(L_o[k]["mean"]), float(v["mean"])
        else:
            a, c = float(L_o[k]), float(v)
        if not fp32 and k.startswith("IoU>="):
            continue                              # threshold counts: discrete in the predictions
        tol = 1e-4 if fp32 else 2e-2
        assert abs(a - c) <= tol * max(1.0, abs(c)), (what, k, a, c)
    report["loss"] = (float(L_o["loss"]), float(L_r["loss"]))
    # |.| / IoU min-max branches may flip for a few sentences under bf16 rounding: looser aggregate bound there
    report["grad_worst"] = check_grads(own, ref, fp32, tol_mult=grad_tol_mult if fp32 else 2.4 * grad_tol_mult, what=what)
    return report
