"""GPU tests of the training step (exoground_b200/train/step.py): the fused flat-buffer AdamW step must track
torch.optim.AdamW with the reference's two-group policy (train/main.py:350-376,418), and a CUDA-graph replay
must equal the eager step."""
import numpy as np
import pytest
import torch

from exoground_b200.model import TemporalAligner
from exoground_b200.model.params import is_no_decay
from exoground_b200.train import loss as exg_loss
from exoground_b200.train.step import TanTrainStep
from exoground_b200.utils import synth

pytestmark = pytest.mark.gpu
E, D, B, T, N, DIN = 2, 3, 4, 32, 8, 64


def _model(dtype):
    torch.manual_seed(3)
    m = TemporalAligner(E, D, use_alignability_head=1, video_embed_dim=DIN, text_embed_dim=DIN, compute_dtype=dtype)
    m.load_state_dict(synth.fill_state_dict({k: v.shape for k, v in m.state_dict().items()}))
    return m


def _batch():
    return synth.tan_batch(B=B, T=T, N=N, d_video=DIN, d_text=DIN, seed=5)


@pytest.mark.parametrize("use_graph", [False, True])
def test_step_matches_torch_adamw_fp32(use_graph):
    args = synth.tan_args(loss_threshold=0.0, use_alignability_head=0)
    b = _batch()
    # eps is raised so that parameters whose gradient is numerical noise (e.g. the key bias, whose true gradient
    # is zero) do not take +-lr sign steps that differ between two correct implementations
    lr, wd, steps, eps = 1e-3, 1e-2, 3, 1e-4
    # ---- ours: fused step
    m = _model(torch.float32)
    m.random_pos_start = 0
    tr = TanTrainStep(m, args, B, T, N, DIN, DIN, lr=lr, wd=wd, eps=eps, use_graph=use_graph)
    tr.load_batch(b)
    losses = [tr.step(b) for _ in range(steps)]
    # ---- same model class driven by torch.optim.AdamW with the reference's policy
    r = _model(torch.float32).cuda()
    r.random_pos_start = 0
    nd = [p for n, p in r.named_parameters() if is_no_decay(n)]
    dc = [p for n, p in r.named_parameters() if not is_no_decay(n)]
    opt = torch.optim.AdamW([{"params": nd, "weight_decay": 0.0}, {"params": dc, "weight_decay": wd}], lr=lr, eps=eps)
    dev = lambda t: t.cuda()
    inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
    ref_losses = []
    for _ in range(steps):
        opt.zero_grad(set_to_none=True)
        out = r(dev(b["video"]), dev(b["text"]), dev(b["video_padding_mask"]), dev(b["text_padding_mask"]), None)
        L = exg_loss.get_loss(inp, dev(b["video"]), dev(b["text"]), dev(b["video_padding_mask"]),
                              dev(b["text_padding_mask"]), out, args, None)
        L["loss"].backward()
        opt.step()
        ref_losses.append(float(L["loss"]))
    assert np.allclose(losses, ref_losses, rtol=1e-4), (losses, ref_losses)
    assert losses[-1] < losses[0]
    pr = dict(r.named_parameters())
    for n, p in m.named_parameters():
        d = (p.detach() - pr[n].detach()).abs().max().item()
        assert d <= 2e-5 + 1e-3 * pr[n].detach().abs().max().item() * lr * steps, (n, d)
    # parameters that never get a gradient are left untouched (torch.optim skips grad=None)
    for n in ("mlp.weight", "text_temporal_pos_embed"):
        assert n in m.flat_params().unused


def test_graph_replay_equals_eager_bf16():
    args = synth.tan_args(loss_threshold=0.5, use_alignability_head=1)
    b = _batch()
    out = []
    for use_graph in (False, True):
        np.random.seed(11)
        m = _model(torch.bfloat16)
        tr = TanTrainStep(m, args, B, T, N, DIN, DIN, lr=1e-3, wd=1e-2, use_graph=use_graph)
        tr.load_batch(b)
        tr.prepare()
        np.random.seed(12)        # identical positional-offset draws for both runs after preparation
        out.append([tr.step(b) for _ in range(3)])
    assert np.allclose(out[0], out[1], rtol=2e-3), out


def test_prefetch_pipeline_matches_plain_steps():
    """step(b, next_batch=nb) (H2D of nb on the copy stream while b computes) == the same steps without prefetch."""
    args = synth.tan_args(loss_threshold=0.0, use_alignability_head=0)
    batches = [synth.tan_batch(B=B, T=T, N=N, d_video=DIN, d_text=DIN, seed=300 + i) for i in range(4)]
    for b in batches:
        for k in ("video", "text", "video_padding_mask", "text_padding_mask"):
            b[k] = b[k].pin_memory()
    runs = []
    for prefetch in (False, True):
        m = _model(torch.bfloat16)
        m.random_pos_start = 0
        # eps raised for the same reason as in test_step_matches_torch_adamw_fp32 (noise-level gradients)
        tr = TanTrainStep(m, args, B, T, N, DIN, DIN, lr=1e-3, eps=1e-4, use_graph=True)
        tr.load_batch(batches[0])
        tr.prepare()
        out = []
        for i, b in enumerate(batches):
            nb = batches[i + 1] if (prefetch and i + 1 < len(batches)) else None
            out.append(tr.step(b, next_batch=nb))
        runs.append(out)
    # fp32 atomics (NCE sums, split-K reduce-add, column sums) make the last bits order-dependent and a bf16
    # rounding flip downstream turns that into ~1e-5..1e-4 on the loss after a few updates (run-to-run, with or
    # without prefetch: scripts/check_determinism.py); a wrong / stale batch would be off by > 1e-2
    assert np.allclose(runs[0], runs[1], rtol=2e-3), runs
    assert min(abs(a - b) for i, a in enumerate(runs[0]) for b in runs[0][i + 1:]) > 1e-2   # the batches really differ


def test_step_async_pipelined_readback_matches_step():
    """step_async(): the loss of step i read after step i+1 has been launched == the blocking step()."""
    args = synth.tan_args(loss_threshold=0.0, use_alignability_head=0)
    batches = [synth.tan_batch(B=B, T=T, N=N, d_video=DIN, d_text=DIN, seed=400 + i) for i in range(5)]
    for b in batches:
        for k in ("video", "text", "video_padding_mask", "text_padding_mask"):
            b[k] = b[k].pin_memory()
    runs = []
    for mode in ("step", "async"):
        m = _model(torch.bfloat16)
        m.random_pos_start = 0
        tr = TanTrainStep(m, args, B, T, N, DIN, DIN, lr=1e-3, eps=1e-4, use_graph=True)
        out, pending = [], None
        for i, b in enumerate(batches):
            nb = batches[i + 1] if i + 1 < len(batches) else None
            if mode == "step":
                out.append(tr.step(b, next_batch=nb))
            else:
                h = tr.step_async(b, next_batch=nb)
                if pending is not None:
                    out.append(pending.result())
                pending = h
        if pending is not None:
            out.append(pending.result())
        runs.append(out)
    assert len(runs[1]) == len(batches)
    assert np.allclose(runs[0], runs[1], rtol=2e-3), runs
    assert min(abs(a - b) for i, a in enumerate(runs[0]) for b in runs[0][i + 1:]) > 1e-2   # the batches really differ


@pytest.mark.parametrize("dtype,use_graph", [(torch.float32, False), (torch.float32, True), (torch.bfloat16, True)])
def test_branch_streams_match_single_stream(dtype, use_graph):
    """The joint-encoder branch on its own CUDA stream (forward and, through autograd's stream replay, backward) must
    give the single-stream step: same losses, same flat gradient after one step, same parameters after three."""
    args = synth.tan_args(loss_threshold=0.5, use_alignability_head=1)
    b = _batch()
    res = []
    for branch in (False, True):
        np.random.seed(11)
        m = _model(dtype)
        tr = TanTrainStep(m, args, B, T, N, DIN, DIN, lr=1e-3, wd=1e-2, eps=1e-4, use_graph=use_graph,
                          branch_streams=branch)
        assert (tr.branch_stream is not None) == branch
        tr.load_batch(b)
        tr.prepare()
        np.random.seed(12)
        losses = [tr.step(b)]
        fp = m.flat_params()
        g1 = fp.grad[:fp.n_used].clone()
        losses += [tr.step(b) for _ in range(2)]
        torch.cuda.synchronize()
        res.append((losses, g1, fp.flat.clone()))
    fp32 = dtype == torch.float32
    assert np.allclose(res[0][0], res[1][0], rtol=1e-5 if fp32 else 2e-3), (res[0][0], res[1][0])
    gn = res[0][1].norm().item()
    assert (res[0][1] - res[1][1]).norm().item() <= (1e-5 if fp32 else 2e-2) * gn
    assert (res[0][2] - res[1][2]).abs().max().item() <= (1e-5 if fp32 else 2e-3)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_exo_step_branch_streams_match_single_stream(dtype):
    """ExoTrainStep with the text unimodal encoder on its own stream == the single-stream step (loss dict, flat
    gradient after one step, parameters after three)."""
    from exoground_b200.model import ExoGroundingTransformer
    from exoground_b200.train.step_exo import ExoTrainStep
    from tests.shapes import exo_shapes
    Dv, Dt, Bx, Tx, Nx = 48, 64, 3, 32, 8
    args = synth.exo_args(use_distill_nce_loss=True, same_view_negative=True)
    batch = synth.exo_batch(B=Bx, T=Tx, N=Nx, d_video=Dv, d_text=Dt, d_ego=Dv, seed=9)
    res = []
    for branch in (False, True):
        m = ExoGroundingTransformer(2, 2, video_embed_dim=Dv, text_embed_dim=Dt, use_distill_nce_loss=True,
                                    random_pos_start=0, compute_dtype=dtype)
        m.load_state_dict(synth.fill_state_dict(exo_shapes(2, 2, Dv, Dt, distill=True)))
        st = ExoTrainStep(m, args, lr=1e-3, wd=1e-2, eps=1e-4, clip_grad=3.0, branch_streams=branch)
        assert (st.branch_stream is not None) == branch
        L = st.step(batch)
        g1 = st.opt.flat_grad().clone()
        losses = [float(L["loss"])] + [float(st.step(batch)["loss"]) for _ in range(2)]
        torch.cuda.synchronize()
        res.append((losses, g1, m.flat_params().flat.clone()))
    fp32 = dtype == torch.float32
    assert np.allclose(res[0][0], res[1][0], rtol=1e-5 if fp32 else 5e-3), (res[0][0], res[1][0])
    assert (res[0][1] - res[1][1]).norm().item() <= (1e-5 if fp32 else 3e-2) * res[0][1].norm().item()
    assert (res[0][2] - res[1][2]).abs().max().item() <= (1e-5 if fp32 else 3e-3)


def test_exo_step_prefetch_matches_plain_steps():
    """step(b, next_batch=nb) (H2D of nb on the copy stream while b computes) == the same steps without prefetch."""
    from exoground_b200.model import ExoGroundingTransformer
    from exoground_b200.train.step_exo import ExoTrainStep
    from tests.shapes import exo_shapes
    Dv, Dt, Bx, Tx, Nx = 48, 64, 3, 32, 8
    args = synth.exo_args(use_distill_nce_loss=True, same_view_negative=True)
    batches = [synth.exo_batch(B=Bx, T=Tx, N=Nx, d_video=Dv, d_text=Dt, d_ego=Dv, seed=20 + i) for i in range(4)]
    batches = [{k: (v.pin_memory() if torch.is_tensor(v) else v) for k, v in b.items()} for b in batches]
    runs = []
    for prefetch in (False, True):
        m = ExoGroundingTransformer(2, 2, video_embed_dim=Dv, text_embed_dim=Dt, use_distill_nce_loss=True,
                                    random_pos_start=0, compute_dtype=torch.float32)
        m.load_state_dict(synth.fill_state_dict(exo_shapes(2, 2, Dv, Dt, distill=True)))
        st = ExoTrainStep(m, args, lr=1e-3, wd=1e-2, eps=1e-4)
        out = []
        for i, b in enumerate(batches):
            nb = batches[i + 1] if (prefetch and i + 1 < len(batches)) else None
            out.append(float(st.step(b, next_batch=nb)["loss"]))
        runs.append(out)
    assert np.allclose(runs[0], runs[1], rtol=1e-4), runs
    assert len({round(x, 3) for x in runs[0]}) == len(runs[0])       # the batches really differ


@pytest.mark.parametrize("use_graph", [False, True])
def test_early_adamw_matches_late(use_graph):
    """AdamW launched from inside backward per finished stack (exg_adamw_dev, device-side lr / bias corrections)
    == the whole AdamW after backward: same parameters after four steps."""
    args = synth.tan_args(loss_threshold=0.5, use_alignability_head=1)
    b = _batch()
    res = []
    for early in (False, True):
        np.random.seed(11)
        m = _model(torch.float32)
        tr = TanTrainStep(m, args, B, T, N, DIN, DIN, lr=1e-3, wd=1e-2, eps=1e-4, use_graph=use_graph,
                          early_adamw=early)
        tr.load_batch(b)
        tr.prepare()
        assert bool(tr._early_ranges) == early
        np.random.seed(12)
        losses = [tr.step(b) for _ in range(4)]
        torch.cuda.synchronize()
        res.append((losses, m.flat_params().flat.clone()))
    # bias corrections: float powf on the host path vs double-then-rounded here -> last-bit differences that four
    # lr = 1e-3 steps amplify to ~1e-5; a wrong step number or a skipped / doubled range would be off by > 1e-2
    assert np.allclose(res[0][0], res[1][0], rtol=1e-4), (res[0][0], res[1][0])
    assert (res[0][1] - res[1][1]).abs().max().item() <= 5e-5


# ---------------------------------------------------------------------- trainer contract (tests/train_contract.py)
from tests import train_contract as tc      # noqa: E402

GRAPH = [False, True]


@pytest.mark.parametrize("use_graph", GRAPH)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_helper_api_after_fused_step(dtype, use_graph):
    tc.check_helpers_after_step("cuda", dtype, use_graph)


@pytest.mark.parametrize("use_graph", GRAPH)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_clip_and_optimizer_state_roundtrip(dtype, use_graph):
    """train/main.py:108-110 (clip_gradients) and :136-146,469 (optimizer state in checkpoints) on the fused step."""
    tc.check_clip_and_state_roundtrip("cuda", dtype, use_graph)


@pytest.mark.parametrize("use_graph", GRAPH)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_cotrain_and_learn_agreement_step(dtype, use_graph):
    """train/main.py:82-89,116-117: online / EMA-target recipe with self-labelling, eager and captured."""
    tc.check_cotrain_step("cuda", dtype, use_graph)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_shape_validation_and_shadow_reload(dtype):
    tc.check_shape_validation_and_shadow_reload("cuda", dtype)


@pytest.mark.parametrize("use_graph", GRAPH)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_repaired_tan_loop(dtype, use_graph, tmp_path):
    """train/main.py:36-157 repaired (exoground_b200/train/tan_loop.py): generic and fused modes agree."""
    tc.check_tan_loop("cuda", dtype, use_graph, tmp_path)


def test_repaired_tan_loop_gradient_accumulation(tmp_path):
    """backprop_freq = 2 (train/main.py:106-113) on the captured step: micro-batch gradients add up in an accumulator
    between optimizer steps."""
    tc.check_tan_loop("cuda", torch.float32, True, tmp_path, backprop_freq=2)


@pytest.mark.parametrize("random_pos", [0, 1])
@pytest.mark.parametrize("distill", [False, True])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_exo_step_graph_matches_eager(dtype, distill, random_pos):
    """ExoTrainStep: CUDA-graph replay == eager device-loss step == eager step through get_loss."""
    tc.check_exo_step_paths("cuda", dtype, distill, random_pos)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_exo_step_gradient_accumulation(dtype):
    """ExoTrainStep(apply_update=False) on the captured step: two micro-batches == one update with the summed gradient."""
    tc.check_exo_accumulation("cuda", dtype)
