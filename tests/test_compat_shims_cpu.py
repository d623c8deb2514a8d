"""SURVEY §7 / §8b "module import names": with exoground_b200/compat first on the module path, the UNMODIFIED
reference trainer's own import lines (`from loss_egoexo4d import get_loss, ...`, `from model.exo_ground_model import
ExoGroundingTransformer`, `from tan_model import TemporalAligner`, `from loss import get_loss`) resolve to the drop-in -
no reference line is edited.  Runs in a subprocess (module names collide with the reference modules other tests load)."""
import os
import subprocess
import sys
import textwrap

import pytest

from oracle import ref_import

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.skipif(not ref_import.available(), reason="reference (tree or oracle/_ref archive) not present")

SCRIPT = textwrap.dedent("""
    import sys, os, types, math, importlib
    ROOT = {root!r}
    sys.path.insert(0, ROOT)
    from oracle import ref_import
    ref_paths = ref_import.paths()              # <ref>/model, <ref>, <ref>/train
    ref_import._install_stubs()
    # search order: repo, compat shims, then the reference tree
    sys.path[:0] = [os.path.join(ROOT, "exoground_b200", "compat")] + ref_paths
    import numpy as np, torch
    from exoground_b200 import ops
    from tests.fake_backend import TorchBackend
    ops.set_backend(TorchBackend())
    import exoground_b200.model as own_model
    import exoground_b200.train.loss_egoexo4d as own_exo_loss
    import exoground_b200.train.loss as own_loss
    with ref_import.quiet():
        trainer = importlib.import_module("main_egoexo4d_distributed")
    assert trainer.ExoGroundingTransformer is own_model.ExoGroundingTransformer
    assert trainer.GroundingModel is own_model.GroundingModel
    assert trainer.ViewInvariantMLP is own_model.ViewInvariantMLP and trainer.ViewInvariantEncoder is own_model.ViewInvariantEncoder
    assert trainer.get_loss is own_exo_loss.get_loss
    assert callable(trainer.visualize) and callable(trainer.save_features_to_dir)
    import model.word2vec_model                  # the namespace overlay still reaches the reference's other modules
    assert "exoground_b200" not in model.word2vec_model.__file__
    import tan_model, loss, tfm_model
    assert tan_model.TemporalAligner is own_model.TemporalAligner and tan_model.TwinTemporalAligner is own_model.TwinTemporalAligner
    assert loss.get_loss is own_loss.get_loss and loss.get_mask_from_time is own_loss.get_mask_from_time
    # the forwarded visualisation entry points reach the reference's own file
    import loss_egoexo4d
    ref_mod = loss_egoexo4d._reference_module()
    assert hasattr(ref_mod, "visualize") and "compat" not in ref_mod.__file__
    # ... and the trainer's own train() runs on what it imported
    from exoground_b200.utils import synth
    class Plot:
        writer = types.SimpleNamespace(flush=lambda: None)
        def add_data(self, *a): pass
        def log_gpustat(self, step=None): pass
    a = synth.exo_args(use_distill_nce_loss=True, same_view_negative=True)
    for k, v in dict(rank=0, views="exo", backprop_freq=1, clip_grad=3.0, iteration=0, prof=None, runtime_save_iter=100,
                     model_path="/tmp", lr=1e-3, wd=1e-2, train_plotter=Plot()).items():
        setattr(a, k, v)
    Dv, Dt, B, T, N = 48, 64, 3, 16, 8
    loader = [synth.exo_batch(B=B, T=T, N=N, d_video=Dv, d_text=Dt, d_ego=Dv, seed=s) for s in (1, 2)]
    for b in loader:
        b["ego_video_features_all"] = b["ego_video_features"]
    m = trainer.ExoGroundingTransformer(2, 2, video_embed_dim=Dv, text_embed_dim=Dt, use_distill_nce_loss=True,
                                        compute_dtype=torch.float32)
    opt = torch.optim.AdamW(trainer.optim_policy(m, a), lr=a.lr, weight_decay=a.wd)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda it: 1.0)
    scaler = torch.amp.GradScaler("cuda", enabled=False)
    with ref_import.quiet():
        avg = trainer.train(loader, m, opt, sched, scaler, torch.device("cpu"), 0, a)
    assert math.isfinite(avg) and a.iteration == 2
    print("COMPAT-OK", avg)
""")


def test_reference_trainer_imports_resolve_to_the_dropin():
    r = subprocess.run([sys.executable, "-c", SCRIPT.format(root=ROOT)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "COMPAT-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
