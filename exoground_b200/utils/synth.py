"""Seeded synthetic inputs of the shapes the reference's loaders emit.

Shapes follow data/loader_egoexo4d_tan.py:333-341 (TAN collate dict: video, text,
padding masks, per-clip start/end second lists) and data/loader_egoexo4d.py:567-598
(ExoGround collate dict).  Everything is generated on the host with explicit
generators so the same call gives the same batch on every box.
"""
from types import SimpleNamespace
import numpy as np
import torch


def tan_batch(B=8, T=64, N=32, d_video=4096, d_text=4096, seed=888, zero_pad_text=True,
              min_video=None, min_text=None, dtype=torch.float32, annot_inside_video=False):
    """One TAN training batch.

    video [B,T,Dv], text [B,N,Dt] ~ N(0,1); video valid length U{T/2..T} (sample 0 full),
    text valid count U{N/2..N} (sample 0 = N, as loss.get_mask_from_time needs max(len)==N,
    train/loss.py:32-42); start ~ sort U(0,T-8), end = start + U(2,8); with annot_inside_video the annotations
    stay inside the un-padded part of each clip (as in real data)."""
    g = torch.Generator().manual_seed(seed)
    rs = np.random.RandomState(seed)
    min_video = T // 2 if min_video is None else min_video
    min_text = N // 2 if min_text is None else min_text
    video = torch.randn(B, T, d_video, generator=g, dtype=torch.float32)
    text = torch.randn(B, N, d_text, generator=g, dtype=torch.float32)
    vlen = rs.randint(min_video, T + 1, size=B); vlen[0] = T
    nlen = rs.randint(min_text, N + 1, size=B); nlen[0] = N
    vmask = torch.zeros(B, T, dtype=torch.bool)
    tmask = torch.zeros(B, N, dtype=torch.bool)
    start, end = [], []
    for b in range(B):
        vmask[b, vlen[b]:] = True
        tmask[b, nlen[b]:] = True
        lim = int(vlen[b]) if annot_inside_video else T
        s = np.sort(rs.uniform(0, max(lim - 8, 1), size=nlen[b])).astype(np.float32)
        e = (s + rs.uniform(2, 8, size=nlen[b])).astype(np.float32)
        if annot_inside_video:
            e = np.minimum(e, np.float32(lim))
        start.append([float(x) for x in s]); end.append([float(x) for x in e])
    if zero_pad_text:
        text[tmask] = 0.0
    return {"video": video.to(dtype), "text": text.to(dtype), "video_padding_mask": vmask,
            "text_padding_mask": tmask, "start": start, "end": end,
            "text_str": [[f"s{b}_{k}" for k in range(int(nlen[b]))] for b in range(B)]}


def tan_args(**kw):
    """The `args` fields train/loss.py::get_loss reads."""
    d = dict(model="init", sim="cos", learn_agreement=0, temporal_agreement_type="keep",
             loss_threshold=0.0, use_alignability_head=0, optim_policy="default", pretrain="")
    d.update(kw)
    return SimpleNamespace(**d)


def exo_batch(B=4, T=64, N=16, d_video=768, d_text=4096, seed=888, V=5, d_ego=None, min_text=4):
    """One ExoGround batch (keys of data/loader_egoexo4d.py:567-598 that the hot path reads)."""
    g = torch.Generator().manual_seed(seed)
    rs = np.random.RandomState(seed)
    video = torch.randn(B, T, d_video, generator=g)
    text = torch.randn(B, N, d_text, generator=g)
    nlen = rs.randint(min_text, N + 1, size=B); nlen[0] = N
    tmask = torch.zeros(B, N, dtype=torch.bool)
    for b in range(B):
        tmask[b, nlen[b]:] = True
    text[tmask] = 0.0
    vmask = torch.zeros(B, T, dtype=torch.bool)
    mean = torch.from_numpy(rs.uniform(0, 1, size=(B, N)).astype(np.float32))
    dur = torch.from_numpy(rs.uniform(0, 0.2, size=(B, N)).astype(np.float32))
    views = ["ego", "best", "2", "3", "worst"]
    per_second_views = [[views[int(i)] for i in rs.randint(0, len(views), size=T)] for _ in range(B)]
    out = {"video_features": video, "narration_features": text, "video_padding_mask": vmask,
           "narration_padding_mask": tmask, "mean": mean, "duration": dur,
           "starts": mean - dur / 2, "ends": mean + dur / 2,
           "metadata": {"per_second_views": per_second_views}}
    if d_ego:
        out["ego_video_features"] = torch.randn(B, V, T, d_ego, generator=g)
        out["view_rank_label"] = torch.from_numpy(rs.randint(0, V, size=(B, T))).long()
        out["view_rank_neg_label"] = torch.from_numpy(rs.randint(0, V, size=(B, T))).long()
        out["same_view_neg_idxs"] = torch.from_numpy(rs.randint(0, T, size=(B, T))).long()
    return out


def exo_args(**kw):
    d = dict(model="joint", use_center_duration=True, iou_loss_eps=1e-8,
             iou_thresholds=[0.1, 0.3, 0.5, 0.7], test="", test_egovlp=False,
             use_distill_nce_loss=False, same_view_negative=False, only_same_view_negative=False)
    d.update(kw)
    return SimpleNamespace(**d)


def fill_state_dict(shapes, seed=888):
    """Deterministic non-degenerate weights for parity tests: {name: shape} -> {name: fp32 tensor}.

    Every tensor has its own generator (crc32(name) ^ seed), so the values do not depend on
    iteration order or on which other parameters exist.  LayerNorm gains are 1+0.1*N(0,1) and all
    biases are non-zero so that no code path is hidden by an identity/zero parameter."""
    import zlib
    out = {}
    for name, shape in shapes.items():
        g = torch.Generator().manual_seed((zlib.crc32(name.encode()) ^ seed) & 0x7FFFFFFF)
        shape = tuple(shape)
        leaf = name.split(".")[-2] if "." in name else name
        if (leaf.startswith("ln_") or ".ln_" in name or name.startswith("ln_")) and name.endswith("weight"):
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif (leaf.startswith("ln_") or name.startswith("ln_")) and name.endswith("bias"):
            t = 0.1 * torch.randn(shape, generator=g)
        elif "pos_embed" in name:
            t = 0.01 * torch.randn(shape, generator=g)
        elif len(shape) == 2:
            t = torch.randn(shape, generator=g) * (shape[1] ** -0.5)
        else:
            t = 0.05 * torch.randn(shape, generator=g)
        out[name] = t
    return out


def htm_align_item(vlen=200, K=40, d_video=1024, d_text=4096, seed=7, aligned_frac=0.3):
    """One HTM-Align-shaped video (eval/eval_zeroshot_align.py:62-77: full-length features, all sentences with
    [aligned, start, end]): video [1, vlen, Dv], sentence embeddings [K, Dt], sorted start seconds, end = start +
    U(2, 8), `aligned` flags (sample 0 / the last sentence are non-alignable so every window has an anchor)."""
    g = torch.Generator().manual_seed(seed)
    rs = np.random.RandomState(seed)
    video = torch.randn(1, vlen, d_video, generator=g)
    text = torch.randn(K, d_text, generator=g)
    start = np.sort(rs.uniform(0, max(vlen - 8, 1), size=K))
    end = np.minimum(start + rs.uniform(2, 8, size=K), vlen - 1)
    aligned = (rs.uniform(size=K) < aligned_frac).astype(np.int64)
    aligned[0] = 0; aligned[-1] = 0
    return {"video": video, "text": text, "start": [float(x) for x in start], "end": [float(x) for x in end],
            "aligned": [int(x) for x in aligned]}
