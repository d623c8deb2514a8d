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


def _run_tan(g, losses_cfg, sub=None):
    E, D, B, T, N, d_in, np_seed = [int(x) for x in g["meta"]]
    m = TemporalAligner(E, D, use_alignability_head=1, video_embed_dim=d_in, text_embed_dim=d_in,
                        compute_dtype=torch.float32)
    m.load_state_dict(synth.fill_state_dict(tan_shapes(E, D, d_in, head=1)))
    b = synth.tan_batch(B=B, T=T, N=N, d_video=d_in, d_text=d_in)
    gen = torch.Generator().manual_seed(1234)
    abs_pos = torch.rand(B, N, 2, generator=gen)
    valid = ~b["text_padding_mask"]
    inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
    for i, (th, al, use_abs) in enumerate(losses_cfg):
        m.zero_grad()
        np.random.seed(np_seed)
        out = m(b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], None)
        args = synth.tan_args(loss_threshold=th, use_alignability_head=al)
        L = exg_loss.get_loss(inp, b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], out,
                              args, abs_pos if use_abs else None)
        for k, v in L.items():
            _close(float(v), g[f"loss{i}/{k}"], rtol=1e-4, atol=1e-5, what=f"loss{i}/{k}")
    L["loss"].backward()
    for k, v in out.items():
        a = v.detach()
        ref = g[f"out/{k}"]
        if k.startswith("logits_"):
            if sub:
                a = a[:, :, ::sub[0], :, ::sub[1]]
            vm = valid[:, ::sub[1]] if sub else valid
            _close(a[:, :, :, vm], ref[:, :, :, vm.numpy()], what=k)
        elif k == "dual_feature_text":
            _close(a[valid], ref[valid.numpy()], what=k)
        elif k == "dual_feature_video" and sub:
            _close(a[:, :, ::sub[0], ::16], ref, atol=5e-5, what=k)
        else:
            _close(a, ref, atol=5e-5, what=k)
    for name, p in m.named_parameters():
        if f"grad/{name}" not in g:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, name
            continue
        gr = p.grad.reshape(-1)
        _close(gr if gr.numel() <= 4096 else gr[::SUB], g[f"grad/{name}"], rtol=2e-3, atol=2e-6, what=f"grad/{name}")


def test_tan_small_host_logic(golden_dir):
    _run_tan(_load(golden_dir, "tan_small"), [(0.0, 0, 0), (0.3, 0, 0), (0.0, 1, 1), (0.5, 1, 0)])


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
