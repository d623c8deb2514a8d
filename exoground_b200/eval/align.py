"""Sliding-window zero-shot alignment inference (reference eval/eval_zeroshot_align.py::test_alignment_htm,
:97-252, 'overlap-seq' method; the caller closure of train/main.py:166-184).

What the reference does per video (vlen seconds, K sentences): windows of `seq_len` seconds every seq_len/4, each
with the sentences whose (non-alignable) ASR timestamps fall near the window; per window one joint + one dual
forward (`get_text_visual_sim_joint/dual`, last stage, / 0.07), added into [K, vlen] accumulators, divided by the
overlap count, averaged (joint + dual) / 2, `sim == 0 -> -6e4`, softmax over time, argmax = predicted second.

B200 design: all windows of a video go through ONE batched forward (batch = #windows, padded to seq_len seconds /
the longest active-sentence run, key-padding masks make padding invisible to the valid positions), the stitch /
overlap average / softmax / argmax stay on the device, and only the [K] predictions are read back.  `batched=False`
runs the reference's window-by-window loop through the same public model API (used by the tests to pin the batched
path).  Sentence embeddings are inputs: the word2vec / BERT text encoders are outside the hot path.
"""
import math

import numpy as np
import torch

from .. import ops
from ..model import functional as Fn

INV_TEMP = 1.0 / 0.07
C = 512


def window_plan(vlen, seq_len, start, end, aligned):
    """The reference's window / active-sentence selection (eval_zeroshot_align.py:128-168) on the host.
    Returns [(step, length, text_left, text_right_inclusive)]."""
    start, end = np.asarray(start, dtype=np.float64), np.asarray(end, dtype=np.float64)
    aligned = np.asarray(aligned).astype(bool)
    K = len(start)
    steps = np.arange(0, vlen - seq_len // 2, seq_len // 4)
    mid = (start + end) / 2
    non_idx = np.arange(K)[~aligned]
    non_mid = mid[~aligned]
    plan = []
    for idx, s in enumerate(steps):
        m = np.logical_and(s - seq_len <= non_mid, non_mid <= s + seq_len + seq_len)
        act = non_idx[m]
        if len(act) == 0:
            continue
        left, right = int(act.min()), int(act.max())
        if idx <= 3:
            left = 0
        elif idx >= len(steps) - 4:
            right = vlen            # the reference's sentinel; slicing clamps it to K - 1
        right = min(right, K - 1)
        if right < left:
            continue
        plan.append((int(s), int(min(vlen, s + seq_len) - s), left, right))
    return plan


def _window_sims_loop(model, video, text, plan, use_head):
    """Reference order: one model call per window (eval_zeroshot_align.py:170-197)."""
    out = []
    for s, ln, l, r in plan:
        v = video[:, s:s + ln]
        t = text[None, l:r + 1]
        sj = model.get_text_visual_sim_joint(v, t)[0, -1].float() * INV_TEMP           # [T, N]
        sd = model.get_text_visual_sim_dual(v, t)[0, -1].float() * INV_TEMP
        ad = aj = None
        if use_head:
            a = model.get_alignability(v, t)
            ad, aj = a['alignability-dual'][0, :, 0].float(), a['alignability-joint'][0, 2, :, 0].float()
        out.append((sj, sd, ad, aj))
    return out


def _window_sims_batched(model, video, text, plan, seq_len, use_head):
    """All windows in one forward: [W, seq_len] seconds x [W, Nmax] sentences with key-padding masks."""
    if model.random_pos_start:
        raise ValueError("batched alignment needs random_pos_start == 0 (evaluation setting)")
    dev = video.device
    W = len(plan)
    T = max(p[1] for p in plan)
    N = max(p[3] - p[2] + 1 for p in plan)
    vlen, K = video.shape[1], text.shape[0]
    st = torch.tensor([p[0] for p in plan], device=dev)
    ln = torch.tensor([p[1] for p in plan], device=dev)
    tl = torch.tensor([p[2] for p in plan], device=dev)
    tn_ = torch.tensor([p[3] - p[2] + 1 for p in plan], device=dev)
    ti = torch.arange(T, device=dev)[None, :]
    ni = torch.arange(N, device=dev)[None, :]
    vmask = ti >= ln[:, None]                                                     # [W, T] True = padding
    tmask = ni >= tn_[:, None]
    vidx = (st[:, None] + ti).clamp(max=vlen - 1)
    tidx = (tl[:, None] + ni).clamp(max=K - 1)
    v = video[0][vidx]                                                           # [W, T, Dv]
    t = text[tidx]                                                               # [W, N, Dt]
    v = v.masked_fill(vmask[:, :, None], 0)
    t = t.masked_fill(tmask[:, :, None], 0)
    wT = model._weights()
    Tc = model.compute_dtype
    E, D = model.num_encoder_layers, model.num_decoder_layers
    vm, tm = vmask.to(torch.uint8).contiguous(), tmask.to(torch.uint8).contiguous()
    xv = model._video_proj(v, wT)
    xt = model._text_proj(t, wT)
    # dual branch (tan_model.py:261-277), last stage only
    st_v = model._video_stages(xv, W, T, vm, 0, None, wT)
    vn = Fn.L2NormFn.apply(st_v[E - 1], Tc).view(W, T, C)
    tnorm = Fn.L2NormFn.apply(xt, Tc).view(W, N, C)
    sd = Fn.SimLogitsFn.apply(vn, tnorm).float() * INV_TEMP                       # [W, T, N]
    # joint branch (tan_model.py:231-258)
    xt_time = model._text_with_time(xt, W, N, 0, None) if model.use_text_pos_enc else xt
    st_j = model._joint_stages(xv, xt_time, W, T, N, vm, tm, 0, None, wT)
    jvn, jtn = Fn.SplitL2NormFn.apply(st_j, W, T, N)
    sj = Fn.SimLogitsFn.apply(jvn[D - 1].view(W, T, C), jtn[D - 1].view(W, N, C)).float() * INV_TEMP
    ad = aj = None
    if use_head:
        ad = model._head(xt, wT).view(W, N).float()
        jt = Fn.GatherRowsFn.apply(st_j, D * W * N, (N, T + N, T), Tc)
        aj = model._head(jt, wT).view(D, W, N)[2].float()
    return sj, sd, ad, aj, (vidx, tidx, vmask, tmask)


@torch.no_grad()
def align_video(model, video, text, start, end, aligned, seq_len=64, use_alignability_head=False, batched=True):
    """video [1, vlen, Dv]; text [K, Dt] sentence embeddings; start / end / aligned: per-sentence lists.
    Returns {'sim' [K, vlen], 'prob' [K, vlen], 'pred' [K] int64 (argmax second), 'align_score' [K]}."""
    dev = video.device
    vlen, K = video.shape[1], text.shape[0]
    plan = window_plan(vlen, seq_len, start, end, aligned)
    a_dual = torch.zeros(K, device=dev)
    a_joint = torch.zeros(K, device=dev)
    tcount = torch.zeros(K, device=dev)
    Kb = ops.get()
    if plan and batched:
        sj, sd, ad, aj, (vidx, tidx, vmask, tmask) = _window_sims_batched(model, video, text, plan, seq_len,
                                                                          use_alignability_head)
        tok = ~tmask
        if use_alignability_head:
            va_d, va_j = ad, aj
        else:                # similarity as alignability (eval_zeroshot_align.py:187-191): max over time
            neg = torch.finfo(torch.float32).min
            va_d = sd.masked_fill(vmask[:, :, None], neg).max(1).values
            va_j = sj.masked_fill(vmask[:, :, None], neg).max(1).values
        a_dual.index_add_(0, tidx[tok], va_d[tok])
        a_joint.index_add_(0, tidx[tok], va_j[tok])
        tcount.index_add_(0, tidx[tok], torch.ones_like(va_d[tok]))
    elif plan:
        T = max(p[1] for p in plan)
        N = max(p[3] - p[2] + 1 for p in plan)
        sj = torch.zeros(len(plan), T, N, device=dev)
        sd = torch.zeros(len(plan), T, N, device=dev)
        for w, ((s, ln, l, r), (wj, wd, ad, aj)) in enumerate(zip(plan, _window_sims_loop(model, video, text, plan,
                                                                                         use_alignability_head))):
            sj[w, :ln, :r - l + 1] = wj
            sd[w, :ln, :r - l + 1] = wd
            if not use_alignability_head:
                ad, aj = wd.max(0).values, wj.max(0).values
            a_dual[l:r + 1] += ad
            a_joint[l:r + 1] += aj
            tcount[l:r + 1] += 1
    if plan:
        # stitch + overlap average + `sim == 0 -> -6e4` + softmax over time + argmax in ONE launch (csrc/align.cu)
        win = torch.tensor(plan, dtype=torch.int32).to(dev)
        sim, prob, pred, row_max = Kb.align_finish(sj.contiguous(), sd.contiguous(), win, K, vlen)
    else:
        sim = torch.full((K, vlen), -6e4, device=dev)
        prob = sim.softmax(-1)
        pred, row_max = prob.argmax(-1), sim.max(-1).values
    eps = torch.tensor(1e-5, device=dev)
    a_dual = a_dual / torch.maximum(tcount, eps)
    a_joint = a_joint / torch.maximum(tcount, eps)
    score = a_joint if use_alignability_head else row_max
    return {'sim': sim, 'prob': prob, 'pred': pred, 'align_score': score, 'windows': len(plan)}


def roc_auc(labels, scores):
    """Rank-based AUC (ties share the mean rank), the quantity sklearn.metrics.roc_auc_score returns."""
    labels, scores = np.asarray(labels).astype(bool), np.asarray(scores, dtype=np.float64)
    order = np.argsort(scores, kind="mergesort")
    ranks = np.empty(len(scores), dtype=np.float64)
    s = scores[order]
    i = 0
    while i < len(s):
        j = i
        while j + 1 < len(s) and s[j + 1] == s[i]:
            j += 1
        ranks[order[i:j + 1]] = (i + j) / 2.0 + 1.0
        i = j + 1
    n_pos, n_neg = labels.sum(), (~labels).sum()
    if n_pos == 0 or n_neg == 0:
        return float('nan')
    return float((ranks[labels].sum() - n_pos * (n_pos + 1) / 2.0) / (n_pos * n_neg))


@torch.no_grad()
def test_alignment(model, dataset, seq_len=64, use_alignability_head=False, batched=True):
    """dataset: iterable of dicts {'video' [1,vlen,Dv], 'text' [K,Dt], 'start', 'end', 'aligned'} (HTM-Align
    shaped).  Returns {'Recall', 'AUC'} as eval_zeroshot_align.py:228-251."""
    recall, all_score, all_tgt = [], [], []
    for item in dataset:
        out = align_video(model, item['video'], item['text'], item['start'], item['end'], item['aligned'], seq_len,
                          use_alignability_head, batched)
        aligned = np.asarray(item['aligned']).astype(bool)
        pred = out['pred'].cpu().numpy()
        s_al = np.asarray(item['start'], dtype=np.float64)[aligned]
        e_al = np.asarray(item['end'], dtype=np.float64)[aligned]
        for p, s, e in zip(pred[aligned], s_al, e_al):
            recall.append(math.floor(s) <= int(p) <= math.ceil(e))
        all_score.append(out['align_score'].float().cpu().numpy())
        all_tgt.append(aligned)
    return {'Recall': float(np.mean(recall)) if recall else float('nan'),
            'AUC': roc_auc(np.concatenate(all_tgt), np.concatenate(all_score))}
