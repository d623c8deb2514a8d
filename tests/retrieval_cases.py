"""Shared body of the clip-retrieval parity test: exoground_b200.eval.retrieval against the
reference's eval/eval_zeroshot_retrieval.py.  The reference's dataset class needs YouCook2 files, so its per-caption
feature recipe (:176-211) is applied here, line by line, to the UNMODIFIED reference model; `compute_metrics` (:13-26)
is the reference's own function, imported."""
import importlib

import numpy as np
import torch

from exoground_b200.eval import retrieval
from oracle import ref_import
from tests import live_ref


def run_retrieval_case(device, dtype, n_items=24, d_in=64, seq_len=16):
    fp32 = dtype == torch.float32
    r, ref, own = live_ref.tan_pair(2, 2, d_in, 0, device, dtype, random_pos_start=0)
    ref.eval(); own.eval()
    with ref_import.quiet():
        ez = importlib.import_module("eval.eval_zeroshot_retrieval")
    g = torch.Generator().manual_seed(5)
    items = []
    for i in range(n_items):
        T = 12 if i % 3 else 24                       # shorter and longer than seq_len (interpolated positions)
        n_clips = 3
        s = [int(x) for x in torch.randint(0, 4, (n_clips,), generator=g)]
        L = 5 + i % 4
        base = torch.randn(1, 1, d_in, generator=g)
        video = (base + 0.5 * torch.randn(n_clips, T, d_in, generator=g)).to(device)
        text = (base[0, 0] + 0.5 * torch.randn(d_in, generator=g)).to(device)
        items.append({"video": video, "text": text, "start_idx": s, "end_idx": [x + L for x in s]})
    # ---- the reference recipe on the reference model
    vs, ts = [], []
    with torch.no_grad():
        for it in items:
            video = it["video"]
            vf = ref.get_visual_feature(video, torch.zeros(video.shape[0:2], device=device, dtype=torch.bool),
                                        interpolate_from=seq_len if video.shape[1] >= seq_len else None)
            vf = vf[:, -1, :]
            tf = ref.get_textual_feature(it["text"].view(1, -1))
            buff = [vf[i, it["start_idx"][i]:it["end_idx"][i], :] for i in range(vf.size(0))]
            vf = torch.stack(buff, 0)
            vf = vf / vf.norm(dim=-1, keepdim=True)
            vf = vf.mean(0).mean(0, keepdim=True)
            vf = vf / vf.norm(dim=-1, keepdim=True)
            tf = tf / tf.norm(dim=-1, keepdim=True)
            vs.append(vf); ts.append(tf)
    V, Tx = torch.cat(vs, 0).cpu().numpy(), torch.cat(ts, 0).cpu().numpy()
    want = {k: float(v) for k, v in ez.compute_metrics(np.dot(Tx, V.T)).items()}
    Vc, Tc = V - V.mean(0, keepdims=True), Tx - Tx.mean(0, keepdims=True)
    wc = ez.compute_metrics(np.dot(Tc, Vc.T))
    ws = ez.compute_metrics(np.dot(Tc / Tc.std(0, keepdims=True), (Vc / Vc.std(0, keepdims=True)).T))
    for k in ("R1", "R5", "R10", "MR"):
        want["C-" + k], want["S-" + k] = float(wc[k]), float(ws[k])
    # ---- the drop-in
    got_v, got_t = [], []
    for it in items:
        v, t = retrieval.segment_features(own, it["video"], it["text"], it["start_idx"], it["end_idx"], seq_len)
        got_v.append(v); got_t.append(t)
    GV, GT = torch.cat(got_v, 0), torch.cat(got_t, 0)
    tol = 1e-4 if fp32 else 2e-2
    assert float((GV.cpu() - torch.as_tensor(V)).abs().max()) <= tol, "visual retrieval features"
    assert float((GT.cpu() - torch.as_tensor(Tx)).abs().max()) <= tol, "text retrieval features"
    got = retrieval.test_retrieval(own, items, seq_len)
    assert set(got) == set(want)
    if fp32:
        for k in want:
            assert abs(got[k] - want[k]) < 1e-9, (k, got[k], want[k])
    else:
        assert abs(got["R1"] - want["R1"]) <= 0.1 and abs(got["MR"] - want["MR"]) <= 1.0, (got, want)
    # the metric function itself against the reference's, on a matrix with ties
    x = np.round(np.random.RandomState(0).randn(40, 40), 1)
    a, b = retrieval.compute_metrics(x), ez.compute_metrics(x)
    assert all(abs(a[k] - float(b[k])) < 1e-12 for k in a)
    return got
