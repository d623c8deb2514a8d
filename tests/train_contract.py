"""Shared bodies of the trainer-contract tests (CPU op-spec backend and CUDA kernels): what the reference's train()
loop (train/main.py:36-157) does around forward/backward - per-parameter clipping, optimizer state in checkpoints,
the 'cotrain' online / EMA-target recipe, the helper API used by in-loop evaluation - on the fused TanTrainStep."""
import copy

import numpy as np
import pytest
import torch

from exoground_b200.model import TemporalAligner, TwinTemporalAligner
from exoground_b200.train import loss as exg_loss
from exoground_b200.train.flat import policy_groups
from exoground_b200.train.step import TanTrainStep
from exoground_b200.utils import synth

E, D, B, T, N, DIN = 2, 3, 4, 32, 8, 64


def _model(dtype, head=1, seed=888):
    m = TemporalAligner(E, D, use_alignability_head=head, video_embed_dim=DIN, text_embed_dim=DIN, compute_dtype=dtype)
    m.load_state_dict(synth.fill_state_dict({k: v.shape for k, v in m.state_dict().items()}, seed=seed))
    return m


def _batch(seed=5, **kw):
    return synth.tan_batch(B=B, T=T, N=N, d_video=DIN, d_text=DIN, seed=seed, **kw)


def _dev(b, device):
    return {k: (v.to(device) if torch.is_tensor(v) else v) for k, v in b.items()}


def _torch_adamw(model, lr, wd, eps):
    nd, dc = policy_groups(model)
    return torch.optim.AdamW([{"params": [p for _, p in nd], "lr": lr, "weight_decay": 0.0},
                              {"params": [p for _, p in dc], "lr": lr, "weight_decay": wd}], lr=lr, weight_decay=wd,
                             eps=eps)


def _manual_step(model, opt, b, args, device, clip=0.0, target=None):
    """One iteration of train/main.py:75-117 with autograd + torch.optim.AdamW on the drop-in model."""
    bd = _dev(b, device)
    inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
    opt.zero_grad(set_to_none=True)
    out = model(bd["video"], bd["text"], bd["video_padding_mask"], bd["text_padding_mask"], None)
    if target is not None:
        with torch.no_grad():
            ema = target(bd["video"], bd["text"], bd["video_padding_mask"], bd["text_padding_mask"], None)
        out = {**out, **{f"ema-{k}": v for k, v in ema.items()}}
    L = exg_loss.get_loss(inp, bd["video"], bd["text"], bd["video_padding_mask"], bd["text_padding_mask"], out, args, None)
    L["loss"].backward()
    if clip > 0:                      # utils/train_utils.py:3-13
        for p in model.parameters():
            if p.grad is not None:
                c = clip / (p.grad.data.norm(2) + 1e-6)
                if c < 1:
                    p.grad.data.mul_(c)
    opt.step()
    return float(L["loss"])


def _assert_params_close(a, b, lr, steps, what, mult=1.0):
    pb = dict(b.named_parameters())
    for n, p in a.named_parameters():
        x, y = p.detach(), pb[n].detach()
        if n.endswith("in_proj_bias"):
            # the key bias has an identically-zero gradient (softmax is shift-invariant per query): what AdamW normalises
            # there is rounding noise, different in any two correct runs -> compare the q and v thirds
            keep = torch.cat((torch.arange(0, 512), torch.arange(1024, 1536))).to(x.device)
            x, y = x[keep], y[keep]
        diff = (x - y).abs()
        d = diff.max().item()
        bound = mult * (2e-5 + 1e-3 * pb[n].detach().abs().max().item() * lr * steps)
        if mult > 1.0:
            # bf16 runs: the kernels' split-K / atomic reductions differ in the last bits between two correct runs, and
            # AdamW turns a rounding-level difference on an element whose gradient is ~0 into up to lr per step on that
            # element: almost every element within the bound, none further apart than the optimizer can move it
            frac = (diff > bound).float().mean().item()
            assert frac <= 2e-3 and d <= 2.1 * lr * steps + 2e-5, (what, n, d, frac)
        else:
            assert d <= bound, (what, n, d)


def check_helpers_after_step(device, dtype, use_graph):
    """after TanTrainStep has installed its device-side positional offsets, the helper API the in-loop
    evaluation calls (eval_zeroshot_align.py:179 -> get_text_visual_sim_*) must keep working and draw on the host."""
    m = _model(dtype)
    tr = TanTrainStep(m, synth.tan_args(), B, T, N, DIN, DIN, device=device, use_graph=use_graph)
    b = _batch()
    tr.step(b)
    bd = _dev(b, device)
    m.eval()
    with torch.no_grad():
        np.random.seed(3)
        a = m.get_text_visual_sim_dual(bd["video"], bd["text"])
        j = m.get_text_visual_sim_joint(bd["video"], bd["text"])
        v = m.get_visual_feature(bd["video"], bd["video_padding_mask"])
        t = m.get_textual_feature_with_time(bd["text"])
        al = m.get_alignability(bd["video"], bd["text"])
        np.random.seed(3)
        a2 = m.get_text_visual_sim_dual(bd["video"], bd["text"])
    assert a.shape == (B, E, T, N) and j.shape == (B, D, T, N) and v.shape == (B, E, T, 512) and t.shape == (B, N, 512)
    assert al["alignability-joint"].shape == (B, D, N, 1)
    assert torch.equal(a, a2)                          # host draw: same numpy seed, same offsets
    m.train()
    tr.step(b)                                         # and the trainer still runs afterwards


def check_clip_and_state_roundtrip(device, dtype, use_graph):
    """clip_grad on the fused path == clip_gradients + torch AdamW; optimizer state written in torch's layout, loaded
    into a fresh fused step AND into a real torch.optim.AdamW, continues identically."""
    fp32 = dtype == torch.float32
    args = synth.tan_args(loss_threshold=0.0, use_alignability_head=0)
    lr, wd, eps, clip = 1e-3, 1e-2, 1e-4, 0.05
    b = _batch()
    kw = dict(device=device, use_graph=use_graph, lr=lr, wd=wd, eps=eps, clip_grad=clip)
    # A: three fused steps in a row
    mA = _model(dtype); mA.random_pos_start = 0
    trA = TanTrainStep(mA, args, B, T, N, DIN, DIN, **kw)
    lossA = [trA.step(b) for _ in range(3)]
    # R: the reference recipe on the same model class (autograd, clip_gradients, torch.optim.AdamW)
    if fp32:
        mR = _model(dtype).to(device); mR.random_pos_start = 0
        optR = _torch_adamw(mR, lr, wd, eps)
        lossR = [_manual_step(mR, optR, b, args, device, clip=clip) for _ in range(3)]
        assert np.allclose(lossA, lossR, rtol=1e-4), (lossA, lossR)
        _assert_params_close(mA, mR, lr, 3, "clip vs reference recipe")
        normsA = trA.seg_sumsq.sqrt()
        assert float((normsA > clip).float().mean()) > 0.2, "clip threshold too loose to test anything"
    # B: two steps, checkpoint, resume in a fresh step object, third step
    mB = _model(dtype); mB.random_pos_start = 0
    trB = TanTrainStep(mB, args, B, T, N, DIN, DIN, **kw)
    for _ in range(2):
        trB.step(b)
    ck = {"state_dict": copy.deepcopy(mB.state_dict()), "optimizer": copy.deepcopy(trB.state_dict())}
    assert set(ck["optimizer"]) == {"state", "param_groups"} and len(ck["optimizer"]["param_groups"]) == 2
    mC = _model(dtype, seed=1); mC.random_pos_start = 0
    trC = TanTrainStep(mC, args, B, T, N, DIN, DIN, **kw)
    trC.load_batch(b); trC.prepare()
    mC.load_state_dict({"module." + k: v for k, v in ck["state_dict"].items()})     # DDP-prefixed, as the trainer saves
    trC.load_state_dict(ck["optimizer"])
    assert trC.step_count == 2
    lossC = trC.step(b)
    assert abs(lossC - lossA[2]) <= (1e-5 if fp32 else 2e-3) * max(1.0, abs(lossA[2])), (lossC, lossA[2])
    _assert_params_close(mC, mA, lr, 3, "resume fused", mult=1.0 if fp32 else 30.0)
    if fp32:
        # T: the same checkpoint resumed by a real torch.optim.AdamW over the reference's parameter groups
        mT = _model(dtype, seed=2).to(device); mT.random_pos_start = 0
        mT.load_state_dict(ck["state_dict"])
        optT = _torch_adamw(mT, lr, wd, eps)
        optT.load_state_dict(ck["optimizer"])
        _manual_step(mT, optT, b, args, device, clip=clip)
        _assert_params_close(mT, mA, lr, 3, "resume torch.optim.AdamW")
        # and back: torch's own state_dict loads into the fused step
        trB.load_state_dict(optT.state_dict())
        assert trB.step_count == 3


def check_cotrain_step(device, dtype, use_graph):
    """args.model == 'cotrain' with learn_agreement (train/main.py:82-89,116-117): target forward under no_grad,
    `ema-` logits drive the self-labelling, _momentum_update after the optimizer - fused step vs the manual loop."""
    fp32 = dtype == torch.float32
    args = synth.tan_args(model="cotrain", learn_agreement=1, temporal_agreement_type="keep")
    lr, wd, eps = 1e-3, 1e-2, 1e-4
    b = _batch(seed=4242, annot_inside_video=True)
    sd = synth.fill_state_dict({k: v.shape for k, v in _model(dtype).state_dict().items()})

    def twin():
        tw = TwinTemporalAligner(0.9, E, D, use_alignability_head=1, video_embed_dim=DIN, text_embed_dim=DIN,
                                 compute_dtype=dtype)
        tw.online.load_state_dict(sd)
        tw.target.load_state_dict({k: v * 0.97 for k, v in sd.items()})      # a target that differs from the online
        tw.online.random_pos_start = 0
        return tw
    twA = twin()
    trA = TanTrainStep(twA, args, B, T, N, DIN, DIN, device=device, use_graph=use_graph, lr=lr, wd=wd, eps=eps)
    lossA = [trA.step(b) for _ in range(2)]
    assert all(np.isfinite(lossA))
    twR = twin().to(device)
    optR = _torch_adamw(twR.online, lr, wd, eps)
    lossR = []
    for _ in range(2):
        lossR.append(_manual_step(twR.online, optR, b, args, device, target=twR.target))
        twR._momentum_update()
    tol = 1e-4 if fp32 else 3e-2
    assert np.allclose(lossA, lossR, rtol=tol), (lossA, lossR)
    if fp32:
        _assert_params_close(twA.online, twR.online, lr, 2, "cotrain online")
        _assert_params_close(twA.target, twR.target, lr, 2, "cotrain target")
    # the target moved (EMA) and stayed gradient-free
    t0 = sd["video_pre_proj.weight"].to(device) * 0.97
    assert not torch.equal(twA.target.video_pre_proj.weight.detach(), t0)
    assert all(p.grad is None for p in twA.target.parameters())
    # learn_agreement on a plain 'init' model through the fused step (needs the video padding mask: ADVICE r1)
    mI = _model(dtype); mI.random_pos_start = 0
    trI = TanTrainStep(mI, synth.tan_args(learn_agreement=1), B, T, N, DIN, DIN, device=device, use_graph=use_graph)
    assert np.isfinite(trI.step(b))
    with pytest.raises(ValueError, match="TwinTemporalAligner"):
        TanTrainStep(_model(dtype), args, B, T, N, DIN, DIN, device=device, use_graph=False)


def check_shape_validation_and_shadow_reload(device, dtype):
    m = _model(dtype)
    tr = TanTrainStep(m, synth.tan_args(), B, T, N, DIN, DIN, device=device, use_graph=False)
    b = _batch()
    tr.step(b)
    bad = synth.tan_batch(B=1, T=T, N=N, d_video=DIN, d_text=DIN, seed=5)
    with pytest.raises(ValueError, match="built for"):
        tr.step(bad)
    if dtype == torch.bfloat16:
        # a checkpoint loaded after the trainer took over the shadow must reach the bf16 weights
        new = synth.fill_state_dict({k: v.shape for k, v in m.state_dict().items()}, seed=77)
        m.load_state_dict(new)
        fp = m.flat_params()
        assert torch.equal(fp.shadow_view("video_pre_proj.weight"), m.video_pre_proj.weight.detach().to(torch.bfloat16))


def check_tan_loop(device, dtype, use_graph, tmp_path, backprop_freq=1):
    """The repaired train/main.py::train loop (exoground_b200/train/tan_loop.py): generic mode (autograd + torch AdamW +
    clip_gradients + LambdaLR + runtime checkpoint) and fused mode (TanTrainStep) over the same synthetic loader in the
    collate format of data/loader_egoexo4d_tan.py:333-341 give the same epoch losses and parameters."""
    import math
    import types
    from exoground_b200.train import tan_loop
    fp32 = dtype == torch.float32
    lr, wd, eps, clip = 1e-3, 1e-2, 1e-4, 0.05

    def loader():
        out = []
        for s in (5, 6, 7):
            b = _batch(seed=s)
            out.append({"video": b["video"], "padding_mask": b["video_padding_mask"].long(), "start": b["start"],
                        "end": b["end"], "narration_features": b["text"],
                        "narration_padding_mask": b["text_padding_mask"], "metadata": {}})
        return out

    def args_(path):
        a = synth.tan_args(loss_threshold=0.0, use_alignability_head=0)
        rows = []
        plot = types.SimpleNamespace(add_data=lambda k, v, it: rows.append((k, float(v), it)))
        for k, v in dict(rank=0, backprop_freq=backprop_freq, clip_grad=clip, iteration=0, runtime_save_iter=2,
                         model_path=str(path), train_plotter=plot).items():
            setattr(a, k, v)
        a.rows = rows
        return a

    def sched(opt_or_none, get_lr=None):
        lam = lambda it: (it + 1) / 2 if it < 2 else 0.5 * (1 + math.cos(math.pi * (it - 2) / 10))
        if opt_or_none is None:      # a scheduler needs an optimizer object: one dummy parameter
            opt_or_none = torch.optim.SGD([torch.nn.Parameter(torch.zeros(1))], lr=lr)
        return torch.optim.lr_scheduler.LambdaLR(opt_or_none, lam)

    import os
    pg, pf = tmp_path / "generic", tmp_path / "fused"
    os.makedirs(pg), os.makedirs(pf)
    # generic
    mG = _model(dtype).to(device); mG.random_pos_start = 0
    optG = _torch_adamw(mG, lr, wd, eps)
    aG = args_(pg)
    scaler = torch.amp.GradScaler("cuda", enabled=False)
    sG = sched(optG)
    avgG = [tan_loop.train(loader(), mG, optG, sG, scaler, device, ep, aG) for ep in range(2)]
    # fused
    mF = _model(dtype); mF.random_pos_start = 0
    aF = args_(pf)
    trF = TanTrainStep(mF, aF, B, T, N, DIN, DIN, device=device, use_graph=use_graph, lr=lr, wd=wd, eps=eps,
                       clip_grad=clip)
    sF = sched(None)
    avgF = [tan_loop.train(loader(), mF, None, sF, None, device, ep, aF, fused_step=trF) for ep in range(2)]
    assert aG.iteration == aF.iteration == 6
    assert np.allclose(avgG, avgF, rtol=1e-4 if fp32 else 3e-2), (avgG, avgF)
    if fp32:
        _assert_params_close(mF, mG, lr, 6, "tan_loop fused vs generic")
    # both wrote a runtime checkpoint whose optimizer state a torch.optim.AdamW over the policy groups accepts
    for path, m in ((pg, mG), (pf, mF)):
        ck = torch.load(os.path.join(path, "runtime.pth.tar"), weights_only=False)
        assert ck["iteration"] == 6 and set(ck["optimizer"]) == {"state", "param_groups"}
        _torch_adamw(m, lr, wd, eps).load_state_dict(ck["optimizer"])
    assert any(k == "train/lr" for k, _, _ in aG.rows) and any(k == "global/loss" for k, _, _ in aF.rows)


def check_exo_step_paths(device, dtype, distill, random_pos=1):
    """ExoTrainStep: the device-loss step (one read-back; captured into a CUDA graph on the GPU) against the eager step
    through the drop-in get_loss (train/main_egoexo4d_distributed.py:105-222): same loss dictionary, same parameters."""
    from exoground_b200.model import ExoGroundingTransformer
    from exoground_b200.train.step_exo import ExoTrainStep
    fp32 = dtype == torch.float32
    Dv, Dt, Bx, Tx, Nx = 48, 64, 3, 16, 8
    args = synth.exo_args(use_distill_nce_loss=distill, same_view_negative=distill)
    batches = [synth.exo_batch(B=Bx, T=Tx, N=Nx, d_video=Dv, d_text=Dt, d_ego=Dv, seed=s) for s in (1, 2, 3)]

    def run(mode):
        m = ExoGroundingTransformer(2, 2, video_embed_dim=Dv, text_embed_dim=Dt, use_distill_nce_loss=distill,
                                    random_pos_start=random_pos, use_text_pos_enc=1, compute_dtype=dtype)
        m.load_state_dict(synth.fill_state_dict({k: v.shape for k, v in m.state_dict().items()}, seed=3))
        st = ExoTrainStep(m, args, lr=1e-3, wd=1e-2, eps=1e-4, clip_grad=0.5, device=device,
                          use_graph=(mode == "graph"))
        if mode == "legacy":
            st.device_loss = False
            st.use_graph = False
        np.random.seed(17)
        out = [st.step(b) for b in batches]
        if mode == "graph":
            assert st.graph is not None
        return m, out
    modes = ["legacy", "device"] + (["graph"] if torch.device(device).type == "cuda" else [])
    res = {md: run(md) for md in modes}
    m0, L0 = res["legacy"]
    for md in modes[1:]:
        m1, L1 = res[md]
        for i, (a, b) in enumerate(zip(L0, L1)):
            assert set(a) == set(b), (md, set(a) ^ set(b))
            # bf16: the first step runs on identical parameters (2e-3); after an AdamW step at lr 1e-3 the two runs'
            # parameters differ by the split-K reduction order of the weight gradients, and the per-rank monitors are
            # means over a handful of seconds (1e-2)
            tol = 1e-5 if fp32 else (3e-3 if i == 0 else 2e-2)
            for k in a:
                va = a[k]["mean"] if isinstance(a[k], dict) else float(a[k])
                vb = b[k]["mean"] if isinstance(b[k], dict) else float(b[k])
                assert abs(va - vb) <= tol * max(1.0, abs(va)), (md, i, k, va, vb)
        _assert_params_close(m1, m0, 1e-3, 3, f"exo step {md}", mult=1.0 if fp32 else 30.0)


def check_exo_accumulation(device, dtype):
    """ExoTrainStep.step(apply_update=False) (`backprop_freq`, main_egoexo4d_distributed.py:175-181): two micro-batches
    accumulated and applied == one update with the SUM of their gradients taken at the same parameters."""
    from exoground_b200.model import ExoGroundingTransformer
    from exoground_b200.train.step_exo import ExoTrainStep
    Dv, Dt, Bx, Tx, Nx = 48, 64, 3, 16, 8
    args = synth.exo_args()
    batches = [synth.exo_batch(B=Bx, T=Tx, N=Nx, d_video=Dv, d_text=Dt, d_ego=Dv, seed=s) for s in (11, 12)]

    def make(use_graph):
        m = ExoGroundingTransformer(2, 2, video_embed_dim=Dv, text_embed_dim=Dt, random_pos_start=0, use_text_pos_enc=1,
                                    compute_dtype=dtype)
        m.load_state_dict(synth.fill_state_dict({k: v.shape for k, v in m.state_dict().items()}, seed=5))
        return m, ExoTrainStep(m, args, lr=1e-3, wd=1e-2, eps=1e-4, clip_grad=0.5, device=device, use_graph=use_graph)
    on_cuda = torch.device(device).type == "cuda"
    mA, stA = make(on_cuda)
    stA.step(batches[0], apply_update=False)
    pA0 = mA.flat_params().flat.clone()
    assert stA.opt.step_count == 0                     # nothing applied yet
    stA.step(batches[1], apply_update=True)
    assert stA.opt.step_count == 1 and not stA.opt._acc_pending
    assert float(stA.opt._gacc.abs().max()) == 0.0
    mB, stB = make(False)
    grads = []
    for b in batches:
        db = stB.to_device(b)
        if not stB.opt.attached:
            for p in mB.parameters():
                p.grad = None
            stB._fwd_bwd(db)
            stB.opt.attach()
        stB.opt.zero_grad()
        stB._fwd_bwd(db)
        grads.append(stB.opt.flat_grad().clone())
    assert torch.equal(pA0[:grads[0].numel()], mB.flat_params().flat[:grads[0].numel()])      # untouched by the first call
    stB.opt.flat_grad().copy_(grads[0] + grads[1])
    stB.opt.reduce_and_step(clip_grad=0.5)
    _assert_params_close(mA, mB, 1e-3, 1, "exo accumulation", mult=1.0 if dtype == torch.float32 else 30.0)
