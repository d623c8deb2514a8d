"""the reference's UNMODIFIED trainer loop as the caller.

`train/main_egoexo4d_distributed.py::train` (:105-222), its `optim_policy` (:544-570), `clip_gradients`
(utils/train_utils.py:3-13), `torch.optim.AdamW`, the warm-up/cosine `LambdaLR` and the runtime-checkpoint save are
imported from /root/reference and driven twice over the same synthetic loader (collate dict of
data/loader_egoexo4d.py:567-598): once with the reference ExoGroundingTransformer + reference get_loss, once with
the drop-in model + drop-in get_loss swapped in (the op-spec backend stands in for libexg.so on this CPU box).
Average loss, every updated parameter and the checkpoint round trip must agree.

Needs the reference tree, so it only runs in the build container (skipped on the GPU box)."""
import copy
import math
import os
import types

import numpy as np
import pytest
import torch

from exoground_b200 import ops
from exoground_b200.model import ExoGroundingTransformer
from exoground_b200.train import loss_egoexo4d as exo_loss
from exoground_b200.utils import synth
from oracle import ref_import
from tests.fake_backend import TorchBackend

pytestmark = pytest.mark.skipif(not ref_import.available(), reason="reference tree not present")


@pytest.fixture(autouse=True)
def fake_backend():
    prev = ops.set_backend(TorchBackend())
    yield
    ops.set_backend(prev)


class _Plotter:
    def __init__(self):
        self.rows = []
        self.writer = types.SimpleNamespace(flush=lambda: None)

    def add_data(self, k, v, it):
        self.rows.append((k, float(v), it))

    def log_gpustat(self, step=None):
        pass


def _args(tmp, **kw):
    a = synth.exo_args(use_distill_nce_loss=True, same_view_negative=True)
    for k, v in dict(rank=0, views="exo", backprop_freq=1, clip_grad=3.0, iteration=0, prof=None,
                     runtime_save_iter=2, model_path=str(tmp), lr=1e-3, wd=1e-2, train_plotter=_Plotter()).items():
        setattr(a, k, v)
    for k, v in kw.items():
        setattr(a, k, v)
    return a


def _drive(trainer, model, get_loss, loader, args):
    """The body of main_egoexo4d_distributed.py::main_worker around train(): optimiser groups, LambdaLR, scaler."""
    trainer.get_loss = get_loss
    params = trainer.optim_policy(model, args)
    opt = torch.optim.AdamW(params, lr=args.lr, weight_decay=args.wd)
    total, warm = 12, 2
    sched = torch.optim.lr_scheduler.LambdaLR(
        opt, lambda it: (it + 1) / warm if it < warm else 0.5 * (1 + math.cos(math.pi * (it - warm) / (total - warm))))
    scaler = torch.amp.GradScaler("cuda", enabled=False)
    np.random.seed(21)
    with ref_import.quiet():
        avg = [trainer.train(loader, model, opt, sched, scaler, torch.device("cpu"), ep, args) for ep in range(2)]
    return avg, opt


def test_unmodified_trainer_loop_drives_the_dropin(tmp_path):
    import importlib
    r = ref_import.load()
    with ref_import.quiet():
        trainer = importlib.import_module("main_egoexo4d_distributed")
    ref_get_loss = trainer.get_loss
    Dv, Dt, B, T, N = 48, 64, 3, 16, 8
    loader = [synth.exo_batch(B=B, T=T, N=N, d_video=Dv, d_text=Dt, d_ego=Dv, seed=s) for s in (1, 2, 3)]
    for b in loader:                       # train() feeds [B,T,D] ego features of the positive view
        b["ego_video_features_all"] = b["ego_video_features"]
    with ref_import.quiet():
        ref = r.exo_ground_model.ExoGroundingTransformer(2, 2, video_embed_dim=Dv, text_embed_dim=Dt,
                                                         use_distill_nce_loss=True)
    sd = synth.fill_state_dict({k: v.shape for k, v in ref.state_dict().items()})
    ref.load_state_dict(sd)
    ours = ExoGroundingTransformer(2, 2, video_embed_dim=Dv, text_embed_dim=Dt, use_distill_nce_loss=True,
                                   compute_dtype=torch.float32)
    ours.load_state_dict({"module." + k: v for k, v in sd.items()})       # DDP-prefixed checkpoint (:716-717)
    try:
        a_ref = _args(tmp_path / "ref")
        a_own = _args(tmp_path / "own")
        os.makedirs(a_ref.model_path), os.makedirs(a_own.model_path)
        avg_ref, _ = _drive(trainer, ref, ref_get_loss, copy.deepcopy(loader), a_ref)
        avg_own, opt = _drive(trainer, ours, exo_loss.get_loss, copy.deepcopy(loader), a_own)
    finally:
        trainer.get_loss = ref_get_loss
    assert a_own.iteration == a_ref.iteration == 6
    np.testing.assert_allclose(avg_own, avg_ref, rtol=2e-4)
    # same keys logged (monitoring dict of loss_egoexo4d.py), same values
    kr = {(k, it): v for k, v, it in a_ref.train_plotter.rows if not k.startswith("device/")}
    ko = {(k, it): v for k, v, it in a_own.train_plotter.rows if not k.startswith("device/")}
    assert kr.keys() == ko.keys()
    for key in kr:
        assert abs(kr[key] - ko[key]) <= 2e-3 * max(1.0, abs(kr[key])), (key, kr[key], ko[key])
    # six AdamW steps later the parameters still agree
    pr = dict(ref.named_parameters())
    worst = {}
    for n, p in ours.named_parameters():
        a, b, p0 = p.detach(), pr[n].detach(), sd[n]
        if n.endswith("in_proj_bias"):
            # the key bias has an identically-zero gradient (softmax is shift-invariant per query): what Adam
            # normalises there is rounding noise, in the reference as much as here -> compare the q and v thirds
            keep = torch.cat((torch.arange(0, 512), torch.arange(1024, 1536)))
            a, b, p0 = a[keep], b[keep], p0[keep]
        upd = (b - p0).norm().item()
        worst[n] = (a - b).norm().item() / (upd + 1e-12) if upd > 0 else 0.0
    assert max(worst.values()) < 2e-2, sorted(worst.items(), key=lambda kv: -kv[1])[:5]
    # the runtime checkpoint written by the reference's own save path loads into a fresh drop-in and back
    import glob
    ck = torch.load(sorted(glob.glob(os.path.join(a_own.model_path, "runtime_*.pth.tar")))[-1], weights_only=False)
    assert ck["iteration"] == 6 and set(ck["state_dict"]) == set(ref.state_dict())
    fresh = ExoGroundingTransformer(2, 2, video_embed_dim=Dv, text_embed_dim=Dt, use_distill_nce_loss=True,
                                    compute_dtype=torch.float32)
    fresh.load_state_dict(ck["state_dict"])
    ref.load_state_dict(ck["state_dict"])
    for (k1, v1), (k2, v2) in zip(fresh.state_dict().items(), ours.state_dict().items()):
        assert k1 == k2 and torch.equal(v1, v2)
    opt.load_state_dict(ck["optimizer"])
