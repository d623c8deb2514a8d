"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

fp32 restatement of the TAN loss (train/loss.py): timestamp -> binary target
(:26-43), masked bidirectional MIL-NCE (:234-278), loss thresholding and the
alignability BCE (:280-360), final combination (:362-376).  The self-labelling
agreement branch (:91-232) lives in agreement_oracle.py.
"""
import torch
import torch.nn.functional as F

NEG = -6e4      # train/loss.py:101,248  (fp16-range sentinel kept for parity)
TEMP = 0.07     # train/loss.py:68-69


def circulant(t, dim):
    """train/loss.py:16-23: out[..., i, ..., j] = t[..., (j - i) mod S, ...]; the new axis j is last.
    Known answer (docstring :19-20): circulant([0,1,2]) -> [[0,1,2],[2,0,1],[1,2,0]]."""
    dim = dim if dim >= 0 else t.dim() + dim
    S = t.shape[dim]
    idx = (torch.arange(S)[None, :] - torch.arange(S)[:, None]) % S
    out = t.movedim(dim, -1)[..., idx]            # ..., i, j
    return out.movedim(-2, dim)


def mask_from_time(start_list, end_list, T, N):
    """get_mask_from_time (train/loss.py:26-43) -> bool [B,N,T]; requires max(len)==N."""
    B = len(start_list)
    st = torch.full((B, N), T + 1e2)
    en = torch.full((B, N), -1e2)
    for b in range(B):
        n = len(start_list[b])
        st[b, :n] = torch.tensor(start_list[b], dtype=torch.float32)
        en[b, :n] = torch.tensor(end_list[b], dtype=torch.float32)
    steps = torch.arange(T, dtype=torch.float32)[None, None, :]
    return (st[:, :, None] <= steps) & (steps < en[:, :, None])


def _lse(x, dim):
    return torch.logsumexp(x, dim=dim)


def milnce_terms(logits, tgt_bnt, text_pad):
    """Core of train/loss.py:234-278 for one logits tensor [B,S,T,B,N] (already / 0.07).

    Returns v [S,B*T], t [S,M], row_has_pos [B*T], col_has_pos [M], valid-column index."""
    B, S, T, _, N = logits.shape
    valid = ~text_pad.bool()                                            # [B,N]
    eye = torch.eye(B, dtype=torch.bool)
    pos = tgt_bnt.permute(0, 2, 1)[:, :, None, :] & eye[:, None, :, None]   # [B,T,B,N]
    pos = pos[:, :, valid].reshape(B * T, -1)                               # [B*T,M]
    lg = logits[:, :, :, valid].permute(1, 0, 2, 3).reshape(S, B * T, -1)   # [S,B*T,M]
    lg_pos = torch.where(pos[None], lg, torch.full_like(lg, NEG))
    v = _lse(lg, -1) - _lse(lg_pos, -1)
    t = _lse(lg, -2) - _lse(lg_pos, -2)
    return v, t, pos.any(-1), pos.any(-2), pos


def tan_loss(logits, start_list, end_list, text_pad, T, N, loss_threshold=0.0,
             use_alignability_head=0, optim_policy="default", abs_text_pos=None,
             tgt_override=None):
    """get_loss for args.model='init', args.sim='cos', learn_agreement=0
    (train/loss.py:57-89,234-376).  `logits` is the model output dict."""
    ld = logits["logits_dual"] / TEMP
    lj = logits["logits_joint"] / TEMP
    B = ld.shape[0]
    tgt = mask_from_time(start_list, end_list, T, N) if tgt_override is None else tgt_override
    out = {}
    v_d, t_d, rowp, colp, pos = milnce_terms(ld, tgt, text_pad)
    v_j, t_j, _, _, _ = milnce_terms(lj, tgt, text_pad)
    loss_dual = (v_d[:, rowp].mean() + t_d[:, colp].mean()) / 2
    loss_joint = (v_j[:, rowp].mean() + t_j[:, colp].mean()) / 2
    out["loss-dual"] = loss_dual.detach()
    out["loss-joint"] = loss_joint.detach()
    valid = ~text_pad.bool()

    if loss_threshold > 0 or use_alignability_head:
        with torch.no_grad():
            # per-text max over time of the last-stage own-video logits (:283-286)
            def own_max(lg):
                diag = torch.stack([lg[b, -1, :, b, :] for b in range(B)], 0)   # [B,T,N]
                return diag.permute(1, 0, 2)[:, valid].max(0).values            # [M]
            md, mj = own_max(ld), own_max(lj)
            zs = lambda m: (m - m.mean(0, keepdim=True)) / m.std(0, keepdim=True)
            metric = -(zs(md) + zs(mj))
            keep = metric <= torch.quantile(metric.float(), loss_threshold, -1, keepdim=True)
            pos_th = pos.clone()
            pos_th[:, ~keep] = False
            rowp_th = pos_th.any(-1)
        if loss_threshold > 0:
            out["loss-dual-all"] = loss_dual.detach()
            out["loss-joint-all"] = loss_joint.detach()
            # NB (:298,304): t-loss is first restricted to columns with positives, then by `keep`,
            # so `keep` must index the compacted [colp] axis exactly as the reference does.
            t_d_c, t_j_c = t_d[:, colp], t_j[:, colp]
            loss_dual_th = (v_d[:, rowp_th].mean() + t_d_c[:, keep].mean()) / 2
            loss_joint_th = (v_j[:, rowp_th].mean() + t_j_c[:, keep].mean()) / 2
            out["loss-dual"] = loss_dual_th.detach()
            out["loss-joint"] = loss_joint_th.detach()
        if use_alignability_head:
            with torch.no_grad():
                lab = torch.full_like(metric, 2.0)
                qd = torch.quantile(md.float(), 0.5, keepdim=True)
                qj = torch.quantile(mj.float(), 0.5, keepdim=True)
                lab[(md > qd) & (mj > qj)] = 1.0
                lab[(md < qd) & (mj < qj)] = 0.0
                if abs_text_pos is not None:
                    c = abs_text_pos[valid, :].mean(-1)
                    lab[(c < 0.2) | (c > 0.8)] = 0.0
            la_j = logits["joint_logits_alignability"][:, 2, :, 0][valid][colp]
            sel = lab != 2
            y = lab[sel]
            pw = torch.ones_like(y) * (1 / y.mean() - 1.0)
            bce_j = F.binary_cross_entropy_with_logits(la_j[sel], y, pos_weight=pw)
            out["loss-joint-bce"] = bce_j.detach()
            out["alignability_top1"] = ((la_j[sel] > 0) == y).float().mean()

    nce_w = 0 if optim_policy == "bce" else 1
    if loss_threshold > 0:
        out["loss-total"] = ((loss_dual + loss_joint) / 2).detach()
        loss = (loss_dual_th + loss_joint_th) / 2
    else:
        loss = (loss_dual + loss_joint) / 2
    if use_alignability_head:
        loss = loss * nce_w + bce_j
    out["loss"] = loss
    return out
