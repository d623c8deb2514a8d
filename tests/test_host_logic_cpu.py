"""Host-logic tests (CPU): the drop-in modules + manual backward chaining are run with the torch
restatement of the op specs (tests/fake_backend.py) and compared with the golden fixtures of the
reference.  This validates buffer plumbing / row maps / gradient routing; the CUDA kernels themselves
are checked against the same op specs in tests/test_kernels_gpu.py."""
import os

import numpy as np
import pytest
import torch

from exoground_b200 import ops
from exoground_b200.model import TemporalAligner
from exoground_b200.train import loss as exg_loss
from exoground_b200.utils import synth
from tests.fake_backend import TorchBackend
from tests.shapes import tan_shapes
from tests.tan_cases import run_tan_case

SUB = 997


@pytest.fixture(autouse=True)
def fake_backend():
    prev = ops.set_backend(TorchBackend())
    yield
    ops.set_backend(prev)


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


def _close(a, b, rtol=1e-4, atol=2e-5, what=""):
    a = torch.as_tensor(np.asarray(a), dtype=torch.float32)
    b = torch.as_tensor(np.asarray(b), dtype=torch.float32)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert torch.allclose(a, b, rtol=rtol, atol=atol), f"{what}: max abs err {(a - b).abs().max().item()}"


def test_tan_small_host_logic(golden_dir):
    run_tan_case(_load(golden_dir, "tan_small"), [(0.0, 0, 0), (0.3, 0, 0), (0.0, 1, 1), (0.5, 1, 0)], "cpu",
                 torch.float32)


def test_tan_sim_helpers(golden_dir):
    g = _load(golden_dir, "tan_sim")
    m = TemporalAligner(2, 2, use_alignability_head=1, video_embed_dim=32, text_embed_dim=32,
                        compute_dtype=torch.float32)
    m.load_state_dict(synth.fill_state_dict(tan_shapes(2, 2, 32, head=1)))
    m.eval()
    gen = torch.Generator().manual_seed(77)
    video = torch.randn(2, 24, 32, generator=gen)
    text = torch.randn(2, 5, 32, generator=gen)
    with torch.no_grad():
        _close(m.get_text_visual_sim_joint(video, text, interpolate_from=16), g["joint_interp16"], what="joint16")
        _close(m.get_text_visual_sim_dual(video, text, interpolate_from=16), g["dual_interp16"], what="dual16")
        m.random_pos_start = 0
        _close(m.get_text_visual_sim_joint(video, text), g["joint_p0"], what="joint_p0")
        _close(m.get_text_visual_sim_dual(video, text), g["dual_p0"], what="dual_p0")
        al = m.get_alignability(video, text)
        _close(al["alignability-dual"], g["align_dual"], atol=5e-5, what="align_dual")
        _close(al["alignability-joint"], g["align_joint"], atol=5e-5, what="align_joint")


def test_state_dict_roundtrip_with_ddp_prefix():
    m = TemporalAligner(1, 1, video_embed_dim=16, text_embed_dim=16, compute_dtype=torch.float32)
    sd = {"module." + k: v.clone() for k, v in m.state_dict().items()}
    m2 = TemporalAligner(1, 1, video_embed_dim=16, text_embed_dim=16, compute_dtype=torch.float32)
    m2.load_state_dict(sd)
    for (k1, v1), (k2, v2) in zip(m.state_dict().items(), m2.state_dict().items()):
        assert k1 == k2 and torch.equal(v1, v2)


# ---------------------------------------------------------------------- ExoGround family
def test_exo_small_host_logic(golden_dir):
    from tests.exo_cases import run_exo_case
    run_exo_case(_load(golden_dir, "exo_small"), "cpu", torch.float32)


def test_keystep_small_host_logic(golden_dir):
    from tests.exo_cases import run_keystep_case
    run_keystep_case(_load(golden_dir, "keystep_small"), _load(golden_dir, "exo_small"), "cpu", torch.float32)


def test_exo_audio_and_pairwise_host_logic(golden_dir):
    from tests.exo_cases import run_audio_pairwise_case
    run_audio_pairwise_case(_load(golden_dir, "exo_audio"), "cpu", torch.float32)


def test_vi_mlp_grads_host_logic():
    from tests.exo_cases import run_vi_mlp_grad_case
    run_vi_mlp_grad_case("cpu", torch.float32)


def test_exo_loss_dispatch_and_errors():
    from exoground_b200.train import loss_egoexo4d as exo_loss
    from exoground_b200.model import ExoGroundingTransformer
    assert exo_loss.get_mode_cam_rank([["a", "b", "b"], ["c"]]) == ["b", "c"]
    tm = torch.tensor([[False, False, True], [False, True, True]])
    assert exo_loss.expand_ranks_with_mask(["x", "y"], tm) == ["x", "x", "y"]
    # no-decoder variant: head on every stage of the joint encoder -> [B, E, N, 2]
    m = ExoGroundingTransformer(2, 1, use_decoder=False, video_embed_dim=16, text_embed_dim=24,
                                compute_dtype=torch.float32)
    b = synth.exo_batch(B=2, T=8, N=4, d_video=16, d_text=24)
    out = m(b["video_features"], b["narration_features"], b["video_padding_mask"], b["narration_padding_mask"])
    assert out["interval_preds"].shape == (2, 2, 4, 2)
    # grounding-only dispatch on the decoder variant
    m = ExoGroundingTransformer(1, 1, video_embed_dim=16, text_embed_dim=24, compute_dtype=torch.float32)
    out = m(b["video_features"], b["narration_features"], b["video_padding_mask"], b["narration_padding_mask"])
    L, iou = exo_loss.get_loss(b, out, b["narration_padding_mask"], synth.exo_args(model="grounding", test=True))
    assert iou.numel() == int((~b["narration_padding_mask"]).sum())
    assert any(k.startswith("Rank ") for k in L)
    L["loss"].backward()
    assert m.grounding_head.weight.grad is not None


def test_tan_self_labelling_agreement_host_logic(golden_dir):
    from tests.tan_cases import run_agree_case
    run_agree_case(_load(golden_dir, "tan_agree"), "cpu", torch.float32)


def test_sliding_window_alignment_host_logic(golden_dir):
    """eval/eval_zeroshot_align.py::test_alignment_htm (reference run on synthetic HTM-Align videos) vs
    exoground_b200.eval.align: per-sentence scores, Recall, AUC; batched windows == window loop."""
    from tests.align_cases import run_align_case
    run_align_case(_load(golden_dir, "tan_align"), "cpu", torch.float32, 2e-4)


def test_twin_momentum_update_after_training_layout_change():
    """TwinTemporalAligner._momentum_update (tan_model.py:334-338) is one kernel over the flat parameter buffers: it
    must stay a per-parameter EMA after a trainer has re-laid-out the ONLINE model (never-used parameters moved to
    the tail of its flat buffer), and the frozen target must keep answering forward_from_ema."""
    from exoground_b200.model.tan_model import TwinTemporalAligner
    from exoground_b200.train.step import TanTrainStep
    tw = TwinTemporalAligner(0.9, 1, 1, video_embed_dim=16, text_embed_dim=16, compute_dtype=torch.float32)
    B, T, N = 2, 8, 4
    st = TanTrainStep(tw.online, synth.tan_args(), B, T, N, 16, 16, device="cpu", use_graph=False, lr=1e-2)
    b = synth.tan_batch(B=B, T=T, N=N, d_video=16, d_text=16, zero_pad_text=False)
    st.step(b)
    assert tw.online.flat_params().unused and not tw.target.flat_params().unused
    before = {n: p.detach().clone() for n, p in tw.target.named_parameters()}
    tw._momentum_update()
    on = dict(tw.online.named_parameters())
    moved = 0
    for n, p in tw.target.named_parameters():
        want = before[n] * 0.9 + on[n].detach() * 0.1
        assert torch.allclose(p, want, atol=1e-6), n
        assert not p.requires_grad
        moved += int((p - before[n]).abs().max() > 0)
    assert moved > 10
    out = tw.forward_from_ema(b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], None)
    assert torch.isfinite(out["logits_joint"]).all()


def test_symmetric_infonce_matches_reference_formula():
    """train/loss_nce.py::get_loss == end2end/main_nce.py:30-48 (restated: two F.cross_entropy calls), values and
    gradients."""
    import torch.nn.functional as F
    from exoground_b200.train import loss_nce
    g = torch.Generator().manual_seed(3)
    v = F.normalize(torch.randn(37, 64, generator=g), dim=-1).requires_grad_(True)
    t = F.normalize(torch.randn(37, 64, generator=g), dim=-1).requires_grad_(True)
    sim = v.matmul(t.t()).div(0.07)
    target = torch.arange(37)
    ref = F.cross_entropy(sim, target) + F.cross_entropy(sim.transpose(0, 1), target)
    gv, gt = torch.autograd.grad(ref, (v, t))
    v2, t2 = v.detach().clone().requires_grad_(True), t.detach().clone().requires_grad_(True)
    L = loss_nce.get_loss(v2, t2, None, synth.tan_args())
    _close(L["loss"].detach(), ref.detach(), what="loss")
    _close(L["loss-per-text"], F.cross_entropy(sim, target).detach(), what="loss-per-text")
    _close(L["top1-per-text"], (sim.argmax(-1) == target).float().mean(), what="top1")
    _close(L["top1-per-video"], (sim.argmax(0) == target).float().mean(), what="top1v")
    L["loss"].backward()
    _close(v2.grad, gv, atol=1e-5, what="dv")
    _close(t2.grad, gt, atol=1e-5, what="dt")


def test_clip_retrieval_head_host_logic():
    """eval/eval_zeroshot_retrieval.py:157-258 (YouCook2 retrieval head) on the op-spec backend."""
    from oracle import ref_import
    if not ref_import.available():
        pytest.skip("reference not present")
    from tests.retrieval_cases import run_retrieval_case
    run_retrieval_case("cpu", torch.float32)


def test_warmup_cosine_schedule_matches_lambda_lr():
    """train/schedule.py against torch's LambdaLR over the reference's lr_schedule_fn (main_egoexo4d_distributed.py:
    832-843, restated here from its definition): same lr at every iteration the loop asks for, resume included."""
    import functools
    import math
    import types
    from exoground_b200.train.schedule import WarmupCosine, lr_multiplier

    def lr_schedule_fn(iteration, iter_per_epoch, args):
        if iteration < args.warmup_iterations:
            return iteration / args.warmup_iterations
        return 0.5 * (1. + math.cos(math.pi * (iteration - args.warmup_iterations)
                                    / (args.epochs * iter_per_epoch - args.warmup_iterations)))
    args = types.SimpleNamespace(warmup_iterations=7, epochs=3)
    ipe, base = 11, 2e-4
    opt = torch.optim.SGD([torch.nn.Parameter(torch.zeros(1))], lr=base)
    ref = torch.optim.lr_scheduler.LambdaLR(opt, functools.partial(lr_schedule_fn, iter_per_epoch=ipe, args=args))
    tgt = types.SimpleNamespace(lr=None)
    own = WarmupCosine(base, ipe, args.epochs, warmup_iterations=7, targets=[tgt])
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        ref.step(0)
        assert own.get_last_lr()[0] == pytest.approx(ref.get_last_lr()[0], abs=1e-15)
        for it in range(1, 3 * ipe + 1):
            ref.step(it)
            own.step(it)
            assert own.get_last_lr()[0] == pytest.approx(ref.get_last_lr()[0], rel=1e-12, abs=1e-18), it
            assert tgt.lr == own.get_last_lr()[0]
    assert lr_multiplier(0, ipe, 3, 7) == 0.0 and lr_multiplier(3 * ipe, ipe, 3, 7) == pytest.approx(0.0, abs=1e-12)
    resumed = WarmupCosine(base, ipe, args.epochs, warmup_iterations=7)
    resumed.load_state_dict(own.state_dict())
    assert resumed.get_last_lr() == own.get_last_lr()
    own.step()                                     # argument-free step advances by one
    assert own.last_epoch == 3 * ipe + 1
