"""Self-labelling agreement targets of the reference's `learn_agreement` branch (train/loss.py:91-232), restated on
the block-diagonal representation used by the CUDA loss: everything here works on the OWN-clip block [B, T, N] of
the last-stage logits (the only entries the reference reads), under no_grad, with device tensor ops on
B*T*N- and B*N*T*T-sized tensors (262 k / 16.8 M elements at B=128, T=64, N=32) — plumbing around the NCE kernels,
not part of their arithmetic.

Per sentence (b, n) with annotated duration d = max(#annotated seconds, 1):
  * p[b, t, n] = softmax_t( softmax_n(logits) / 0.07 )                                  (loss.py:108-110, 160-162)
  * every window [s, s + d) that fits inside the clip is scored by the mean of p over its seconds, seconds 0 and
    T-1 never counting (loss.py:117-139); the best window is the sentence's self-label, the mean logit over it its
    confidence (loss.py:141-150)
  * joint and dual self-labels are compared (IoU over seconds, 0.3-quantile confidence, loss.py:181-201) and merged
    into the annotation according to `temporal_agreement_type` (loss.py:203-216)
  * exclusion principle: at each second only the first sentence keeps a 1; sentences left with nothing get their
    annotation back (loss.py:218-228).
"""
import torch

NEG = -6e4


def own_block(logits_sm, stage, B, T, N, clip0=0):
    """[B, T, N] fp32 = logits[stage, (b, t), (b + clip0, n)] of stage-major logits [S, B*T, Bc*N]."""
    blk = logits_sm[stage].view(B, T, -1, N)[:, :, clip0:clip0 + B]
    return torch.diagonal(blk, dim1=0, dim2=2).permute(2, 0, 1).float()


def _windows(dur, T):
    """w[b, n, s, t]: averaging weights of the window of sentence (b, n) starting at second s (zero rows for windows
    that do not fit; seconds 0 and T-1 carry no weight)."""
    dev = dur.device
    s = torch.arange(T, device=dev).view(1, 1, T, 1)
    t = torch.arange(T, device=dev).view(1, 1, 1, T)
    d = dur.view(*dur.shape, 1, 1)
    inside = (t >= s) & (t < s + d) & (s + d <= T) & (d > 0)
    inside = inside & (t != 0) & (t != T - 1)
    cnt = inside.sum(-1, keepdim=True).float().clamp(min=1e-3)
    return inside.float() / cnt


def _best_window(lg, w):
    """lg [B, T, N] masked own logits, w [B, N, T, T] -> (self label [B, N, T] bool, mean logit over it [B, N])."""
    prob = (lg.softmax(-1) / 0.07).softmax(-2)                     # [B, T, N]
    p_bnt = prob.permute(0, 2, 1)
    scan = (p_bnt[:, :, None, :] * w).sum(-1)                       # [B, N, T(start)]
    pos = scan.max(-1).indices
    best = torch.gather(w, 2, pos[:, :, None, None].expand(-1, -1, 1, w.shape[-1])).squeeze(2)   # [B, N, T]
    conf = (lg.permute(0, 2, 1) * best).sum(-1)
    return best > 0, conf


def _quantile(x, mask, q):
    from .loss import _masked_quantile          # sort + gather: no host synchronisation (CUDA-graph safe)
    return _masked_quantile(x.reshape(-1).float(), mask.reshape(-1), q)


def self_label_targets(own_joint, own_dual, tgt_raw, video_pad, text_pad, mode):
    """own_*: [B, T, N] own-clip last-stage logits already divided by 0.07; tgt_raw bool [B, N, T] (annotation);
    video_pad bool [B, T]; text_pad bool [B, N]; mode in {'i', 'u', 'keep', 'keep-joint'}.
    Returns (targets bool [B, N, T], {'confidence-ratio', 'iou-threshold'})."""
    B, T, N = own_joint.shape
    dev = own_joint.device
    vp, tp = video_pad.bool(), text_pad.bool()
    drop = vp[:, :, None] | tp[:, None, :]
    lj = own_joint.masked_fill(drop, NEG)
    ld = own_dual.masked_fill(drop, NEG)
    dur = tgt_raw.sum(-1).clamp(min=1).masked_fill(tp, 0)           # [B, N]
    w = _windows(dur, T)
    j_tgt, j_conf = _best_window(lj, w)
    d_tgt, d_conf = _best_window(ld, w)

    inter, union = j_tgt & d_tgt, j_tgt | d_tgt
    iou = inter.sum(-1).float() / union.sum(-1).float().clamp(min=1e-5)               # [B, N]
    valid = ~tp
    conf_text = (d_conf >= _quantile(d_conf, valid, 0.3)) & (j_conf >= _quantile(j_conf, valid, 0.3))
    iou_th = torch.full((), 0.5, device=dev)         # (torch.tensor(..) is a host copy: not graph-capturable)
    conf_iou = iou >= iou_th
    conf = conf_text & conf_iou
    if mode == 'i':
        agree = inter & conf[:, :, None]
    elif mode == 'u':
        agree = union & conf[:, :, None]
    elif mode == 'keep':
        agree = torch.where(conf_iou[:, :, None], union, tgt_raw.bool())
    elif mode == 'keep-joint':
        agree = torch.where(conf_iou[:, :, None], j_tgt, tgt_raw.bool())
    else:
        raise ValueError(f"temporal_agreement_type {mode!r}")

    # exclusion principle (loss.py:218-228): per (clip, second) only the first sentence with a 1 keeps it; the
    # reference's scatter also writes sentence 0 where no sentence has a 1 and then restores sentence 0's row.
    a = agree.float()                                               # [B, N, T]
    first = a.argmax(dim=1, keepdim=True)                           # [B, 1, T]
    dedup = torch.zeros_like(a).scatter_(1, first, 1.0)
    dedup[:, 0, :] = a[:, 0, :]
    empty = dedup.sum(-1) == 0                                      # [B, N]
    dedup = torch.where(empty[:, :, None], tgt_raw.float(), dedup)
    stats = {'confidence-ratio': (conf & valid).float().sum() / valid.float().sum(), 'iou-threshold': iou_th}
    return dedup > 0, stats
