"""Edge cases of the TAN path against the UNMODIFIED reference run live (build container only): ragged clips down to
one valid second / one valid sentence, a single-clip batch, sizes that are not multiples of any tile, annotations
that fall outside the clip (rows / columns without a positive), thresholded + alignability losses.  The drop-in runs
on the op-spec backend (CPU); the same cases run through the CUDA kernels in tests/test_model_gpu.py via goldens."""
import numpy as np
import pytest
import torch

from exoground_b200 import ops
from exoground_b200.model import TemporalAligner
from exoground_b200.train import loss as exg_loss
from exoground_b200.utils import synth
from oracle import ref_import
from tests.fake_backend import TorchBackend
from tests.tan_cases import rel_l2

pytestmark = pytest.mark.skipif(not ref_import.available(), reason="reference tree not present")


@pytest.fixture(autouse=True)
def fake_backend():
    prev = ops.set_backend(TorchBackend())
    yield
    ops.set_backend(prev)


def _pair(E, D, d_in, head):
    r = ref_import.load()
    with ref_import.quiet():
        ref = r.tan_model.TemporalAligner(E, D, use_alignability_head=head)
        ref.video_pre_proj = torch.nn.Linear(d_in, 512, bias=False)
        ref.text_pre_proj = torch.nn.Linear(d_in, 512, bias=False)
    sd = synth.fill_state_dict({k: v.shape for k, v in ref.state_dict().items()}, seed=31)
    ref.load_state_dict(sd)
    own = TemporalAligner(E, D, use_alignability_head=head, video_embed_dim=d_in, text_embed_dim=d_in,
                          compute_dtype=torch.float32)
    own.load_state_dict(sd)
    return r, ref, own


CASES = [
    # B, T, N, min_video, min_text, threshold, head, seed
    (4, 12, 6, 1, 1, 0.0, 0, 1),         # clips with a single valid second / sentence
    (1, 9, 5, 9, 5, 0.0, 0, 2),          # one clip: every negative comes from the clip itself
    (3, 20, 7, 2, 2, 0.5, 1, 3),         # odd sizes + quantile threshold + alignability BCE
    (5, 16, 8, 3, 1, 0.3, 0, 4),
]


@pytest.mark.parametrize("B,T,N,mv,mt,th,head,seed", CASES)
def test_ragged_and_degenerate_batches_match_reference(B, T, N, mv, mt, th, head, seed):
    d_in = 24
    r, ref, own = _pair(2, 3, d_in, head)          # the alignability BCE reads joint stage 3 (train/loss.py:344)
    b = synth.tan_batch(B=B, T=T, N=N, d_video=d_in, d_text=d_in, seed=seed, zero_pad_text=False, min_video=mv,
                        min_text=mt)
    # push an annotation past the end of the clip: rows and a column without any positive.  (Not with the threshold:
    # the reference then indexes a vector over the columns WITH positives by a mask over the VALID columns and
    # raises IndexError, train/loss.py:299.)
    if th == 0:
        b["start"][-1][-1], b["end"][-1][-1] = float(T + 3), float(T + 5)
    inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
    args = synth.tan_args(loss_threshold=th, use_alignability_head=head)
    outs = []
    for mod, get_loss in ((ref, r.loss.get_loss), (own, exg_loss.get_loss)):
        np.random.seed(77)
        with ref_import.quiet():
            out = mod(b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], None)
            L = get_loss(inp, b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], out, args, None)
        L["loss"].backward()
        outs.append((out, L, {n: p.grad for n, p in mod.named_parameters() if p.grad is not None}))
    (o_r, L_r, g_r), (o_o, L_o, g_o) = outs
    valid = ~b["text_padding_mask"]
    for k in ("logits_dual", "logits_joint"):
        a, c = o_o[k].detach()[:, :, :, valid], o_r[k].detach()[:, :, :, valid]      # padded columns: reference trap 2
        assert torch.allclose(a, c, rtol=1e-4, atol=2e-5), (k, (a - c).abs().max().item())
    assert set(L_o) == set(L_r)
    for k in L_r:
        a, c = float(L_o[k]), float(L_r[k])
        assert abs(a - c) <= 1e-4 * max(1.0, abs(c)), (k, a, c)
    assert set(g_o) == set(g_r)
    worst = max((rel_l2(g_o[n], g_r[n]), n) for n in g_r if g_r[n].norm() > 1e-7 and not n.endswith("in_proj_bias"))
    assert worst[0] <= 2e-3, worst
    # predicted timestamps: argmax over time of the last-stage (joint + dual) / 2 logits, bit-exact
    for bi in range(B):
        s_o = (o_o["logits_dual"][bi, -1, :, bi] + o_o["logits_joint"][bi, -1, :, bi]) / 2
        s_r = (o_r["logits_dual"][bi, -1, :, bi] + o_r["logits_joint"][bi, -1, :, bi]) / 2
        keep = valid[bi]
        assert torch.equal(s_o[:, keep].argmax(0), s_r[:, keep].argmax(0))


EXO_CASES = [
    # B, T, N, min_text, use_center_duration, same_view_negative, only_same, decoder, text_pos, seed
    (3, 10, 5, 1, True, True, False, True, 0, 1),       # clips down to ONE narration, odd T / N
    (1, 7, 3, 3, False, False, False, True, 1, 2),      # a single clip, start / end parametrisation, text pos-enc
    (2, 12, 6, 2, True, True, True, False, 0, 3),       # no decoder: head on every joint stage; only-same-view negatives
]


@pytest.mark.parametrize("B,T,N,mt,cd,svn,only,dec,tpos,seed", EXO_CASES)
def test_exoground_ragged_batches_match_reference(B, T, N, mt, cd, svn, only, dec, tpos, seed):
    """ExoGroundingTransformer + loss_egoexo4d.get_loss + backward against the reference run live on ragged batches."""
    from exoground_b200.model import ExoGroundingTransformer
    from exoground_b200.train import loss_egoexo4d as exo_loss
    Dv, Dt = 24, 40
    r = ref_import.load()
    with ref_import.quiet():
        ref = r.exo_ground_model.ExoGroundingTransformer(2, 2, use_decoder=dec, video_embed_dim=Dv, text_embed_dim=Dt,
                                                         use_distill_nce_loss=True, use_text_pos_enc=tpos)
    sd = synth.fill_state_dict({k: v.shape for k, v in ref.state_dict().items()}, seed=41)
    ref.load_state_dict(sd)
    own = ExoGroundingTransformer(2, 2, use_decoder=dec, video_embed_dim=Dv, text_embed_dim=Dt,
                                  use_distill_nce_loss=True, use_text_pos_enc=tpos, compute_dtype=torch.float32)
    own.load_state_dict(sd)
    b = synth.exo_batch(B=B, T=T, N=N, d_video=Dv, d_text=Dt, d_ego=Dv, seed=seed, min_text=mt)
    args = synth.exo_args(use_distill_nce_loss=True, use_center_duration=cd, same_view_negative=svn,
                          only_same_view_negative=only)
    outs = []
    for mod, get_loss in ((ref, r.loss_egoexo4d.get_loss), (own, exo_loss.get_loss)):
        np.random.seed(5)
        with ref_import.quiet():
            out = mod(b["video_features"], b["narration_features"], b["video_padding_mask"],
                      b["narration_padding_mask"], egocentric_video_embed=b["ego_video_features"][:, 0])
            if not dec:          # [B,E,N,2] without the decoder: the loss reads the last stage
                out = dict(out, interval_preds=out["interval_preds"][:, -1])
            L, iou = get_loss(b, out, b["narration_padding_mask"], args)
        L["loss"].backward()
        outs.append((out, L, iou, {n: p.grad for n, p in mod.named_parameters() if p.grad is not None}))
    (o_r, L_r, i_r, g_r), (o_o, L_o, i_o, g_o) = outs
    for k in o_r:
        a, c = o_o[k].detach(), o_r[k].detach()
        assert a.shape == c.shape and torch.allclose(a, c, rtol=1e-4, atol=3e-5), (k, (a - c).abs().max().item())
    assert torch.allclose(i_o.detach().cpu(), i_r.detach().cpu(), rtol=1e-4, atol=1e-5)
    for k, v in L_r.items():
        if isinstance(v, dict):          # per-camera-rank breakdown {'mean': ...}
            assert abs(float(L_o[k]["mean"]) - float(v["mean"])) <= 1e-4 * max(1.0, abs(float(v["mean"]))), k
        else:
            assert abs(float(L_o[k]) - float(v)) <= 1e-4 * max(1.0, abs(float(v))), (k, float(L_o[k]), float(v))
    assert set(g_o) == set(g_r)
    worst = max((rel_l2(g_o[n], g_r[n]), n) for n in g_r if g_r[n].norm() > 1e-7 and not n.endswith("in_proj_bias"))
    assert worst[0] <= 2e-3, worst
