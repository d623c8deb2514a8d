"""YouCook2-style zero-shot clip retrieval head: the device part of the reference's
eval/eval_zeroshot_retrieval.py::test_retrieval_yc2 (:157-258) on the drop-in model's kernels.

Per caption: `num_clips` temporal crops of one video segment go through `get_visual_feature` in ONE batched call
(interpolated position table when the crop is longer than the training length, :178-182), the last encoder stage is cut
to the annotated span of each crop, L2-normalised per second, averaged over time and crops and normalised again
(:196-207); the caption embedding goes through `get_textual_feature` and is normalised.  Retrieval = one similarity GEMM
over all (caption, segment) pairs + the rank metrics of MIL-NCE's `compute_metrics` (:13-26), raw / centred /
standardised (:225-256).  Dataset I/O (YouCook2_Feature, tokenizer, word2vec) stays the reference's.
"""
import numpy as np
import torch

from ..model import functional as Fn

C = 512


@torch.no_grad()
def segment_features(model, video, lang_embed, start_idx, end_idx, seq_len=64, sim='cos'):
    """video [num_clips, T, Dv] crops of one segment; lang_embed [1, Dt] caption embedding; start_idx / end_idx
    [num_clips] the annotated span inside each crop.  Returns (visual [1, C], text [1, C]) as the reference builds
    them (eval_zeroshot_retrieval.py:176-211)."""
    n, T, _ = video.shape
    vmask = torch.zeros(n, T, dtype=torch.bool, device=video.device)
    vf = model.get_visual_feature(video, vmask, interpolate_from=seq_len if T >= seq_len else None)   # [n,E,T,C]
    if vf.dim() == 4:
        vf = vf[:, -1]
    vf = vf.float()
    lens = {int(e) - int(s) for s, e in zip(start_idx, end_idx)}
    if len(lens) != 1:
        raise RuntimeError("segment_features: the crops of one segment must keep the same number of seconds "
                           "(torch.stack in the reference fails otherwise, eval_zeroshot_retrieval.py:199)")
    span = torch.stack([vf[i, int(s):int(e)] for i, (s, e) in enumerate(zip(start_idx, end_idx))], 0)  # [n, L, C]
    if sim == 'cos':
        span = span / span.norm(dim=-1, keepdim=True)
    v = span.mean(0).mean(0, keepdim=True)
    t = model.get_textual_feature(lang_embed.view(1, 1, -1)).view(1, C).float()
    if sim == 'cos':
        v = v / v.norm(dim=-1, keepdim=True)
        t = t / t.norm(dim=-1, keepdim=True)
    return v, t


def compute_metrics(x):
    """Rank metrics of a [caption, segment] similarity matrix whose diagonal holds the matching pairs
    (eval_zeroshot_retrieval.py:13-26, from MIL-NCE): R@1/5/10 and the median rank."""
    x = np.asarray(x, dtype=np.float64)
    order = np.sort(-x, axis=1)
    d = -np.diag(x)[:, None]
    ind = np.where(order - d == 0)[1]
    return {'R1': float(np.sum(ind == 0)) / len(ind), 'R5': float(np.sum(ind < 5)) / len(ind),
            'R10': float(np.sum(ind < 10)) / len(ind), 'MR': float(np.median(ind) + 1)}


@torch.no_grad()
def retrieval_metrics(text_features, visual_features, compute_dtype=torch.float32):
    """text_features / visual_features [M, C] (row i of each belongs to pair i).  The [M, M] similarity matrices
    (raw, centred, standardised: eval_zeroshot_retrieval.py:225-256) are one libexg.so GEMM each."""
    t, v = text_features.float(), visual_features.float()

    def sim(a, b):
        return Fn.SimLogitsFn.apply(a.to(compute_dtype).unsqueeze(0).contiguous(),
                                    b.to(compute_dtype).unsqueeze(0).contiguous())[0].float().cpu().numpy()
    m = compute_metrics(sim(t, v))
    tc, vc = t - t.mean(0, keepdim=True), v - v.mean(0, keepdim=True)
    mc = compute_metrics(sim(tc, vc))
    ts, vs = tc / tc.std(0, keepdim=True, unbiased=False), vc / vc.std(0, keepdim=True, unbiased=False)
    ms = compute_metrics(sim(ts, vs))
    for k in ('R1', 'R5', 'R10', 'MR'):
        m['C-' + k] = mc[k]
        m['S-' + k] = ms[k]
    return m


@torch.no_grad()
def test_retrieval(model, items, seq_len=64, sim='cos'):
    """items: iterable of {'video' [num_clips, T, Dv], 'text' [Dt], 'start_idx', 'end_idx'} (what YouCook2_Feature +
    the language model produce per caption).  Returns the metric dict of test_retrieval_yc2."""
    vs, ts = [], []
    for it in items:
        v, t = segment_features(model, it['video'], it['text'], it['start_idx'], it['end_idx'], seq_len, sim)
        vs.append(v)
        ts.append(t)
    return retrieval_metrics(torch.cat(ts, 0), torch.cat(vs, 0),
                             compute_dtype=torch.float32)
