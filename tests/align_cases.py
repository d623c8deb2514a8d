"""Shared body of the sliding-window alignment parity test (CPU op-spec backend and GPU kernels): rebuild the
golden's synthetic HTM-Align videos, run exoground_b200.eval.align and compare with what the UNMODIFIED
eval/eval_zeroshot_align.py::test_alignment_htm produced (tests/golden/tan_align.npz, oracle/make_golden.py)."""
import numpy as np
import torch

from exoground_b200.eval import align
from exoground_b200.model import TemporalAligner
from exoground_b200.utils import synth
from tests.shapes import tan_shapes


def run_align_case(g, device, dtype, score_tol):
    E, D, d_in = (int(x) for x in g["dims"])
    seq_len = int(g["seq_len"])
    m = TemporalAligner(E, D, use_alignability_head=0, video_embed_dim=d_in, text_embed_dim=d_in, random_pos_start=0,
                        compute_dtype=dtype)
    m.load_state_dict(synth.fill_state_dict(tan_shapes(E, D, d_in, head=0)))
    m = m.to(device).eval()
    items = []
    for vlen, K, seed in g["recipe"]:
        it = synth.htm_align_item(vlen=int(vlen), K=int(K), d_video=d_in, d_text=d_in, seed=int(seed))
        it["video"], it["text"] = it["video"].to(device), it["text"].to(device)
        items.append(it)
    res = {}
    for batched in (False, True):
        outs = [align.align_video(m, it["video"], it["text"], it["start"], it["end"], it["aligned"], seq_len,
                                  False, batched) for it in items]
        score = np.concatenate([o["align_score"].float().cpu().numpy() for o in outs])
        ref = g["score"]
        err = np.abs(score - ref).max() / max(np.abs(ref).max(), 1e-6)
        assert err <= score_tol, f"batched={batched}: per-sentence alignment score rel err {err:.3e} > {score_tol}"
        res[batched] = outs
        metrics = align.test_alignment(m, items, seq_len, False, batched)
        if dtype == torch.float32:       # argmax-derived metrics: exact in fp32 (north star: indices bit-exact)
            assert abs(metrics["Recall"] - float(g["recall"])) < 1e-12, (metrics, float(g["recall"]))
            assert abs(metrics["AUC"] - float(g["auc"])) < 1e-9, (metrics, float(g["auc"]))
    # one batched forward per video == the reference's window-by-window loop
    for a, b in zip(res[False], res[True]):
        scale = float(a["sim"][a["sim"] > -1e4].abs().max())
        d = (a["sim"] - b["sim"]).abs().max().item() / scale
        assert d <= score_tol, f"batched vs loop sim: {d:.3e}"
        if dtype == torch.float32:
            assert torch.equal(a["pred"], b["pred"])
    assert abs(align.roc_auc(g["tgt"], g["score"]) - float(g["auc"])) < 1e-12     # own AUC == sklearn's
