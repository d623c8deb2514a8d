"""Trainer contract on the op-spec backend (CPU, host logic): see tests/train_contract.py; the same bodies run on the
CUDA kernels (eager and CUDA-graph, fp32 and bf16) in tests/test_train_step_gpu.py."""
import pytest
import torch

from exoground_b200 import ops
from tests import train_contract as tc
from tests.fake_backend import TorchBackend


@pytest.fixture(autouse=True)
def fake_backend():
    prev = ops.set_backend(TorchBackend())
    yield
    ops.set_backend(prev)


def test_helper_api_after_fused_step():
    tc.check_helpers_after_step("cpu", torch.float32, False)


def test_clip_and_optimizer_state_roundtrip():
    tc.check_clip_and_state_roundtrip("cpu", torch.float32, False)


def test_cotrain_and_learn_agreement_step():
    tc.check_cotrain_step("cpu", torch.float32, False)


def test_shape_validation():
    tc.check_shape_validation_and_shadow_reload("cpu", torch.float32)


@pytest.mark.parametrize("backprop_freq", [1, 2])
def test_repaired_tan_loop(tmp_path, backprop_freq):
    """backprop_freq=2: the optimizer steps on iterations 0, 2 of each epoch (train/main.py:107) and the gradients of
    the iterations in between add up, un-normalised - generic loop and fused step (run(apply_update=False)) alike."""
    tc.check_tan_loop("cpu", torch.float32, False, tmp_path, backprop_freq=backprop_freq)


@pytest.mark.parametrize("distill", [False, True])
def test_exo_step_device_loss_matches_get_loss(distill):
    tc.check_exo_step_paths("cpu", torch.float32, distill)


def test_exo_step_gradient_accumulation():
    tc.check_exo_accumulation("cpu", torch.float32)


def test_step_async_returns_the_same_loss_on_the_spec_backend():
    from exoground_b200.train.step import TanTrainStep
    from exoground_b200.utils import synth as sy
    args = sy.tan_args(loss_threshold=0.0, use_alignability_head=0)
    b = tc._batch()
    losses = []
    for mode in ("step", "async"):
        m = tc._model(torch.float32)
        m.random_pos_start = 0
        tr = TanTrainStep(m, args, tc.B, tc.T, tc.N, tc.DIN, tc.DIN, device="cpu", use_graph=False)
        losses.append(tr.step(b) if mode == "step" else tr.step_async(b).result())
    assert losses[0] == pytest.approx(losses[1], rel=1e-6)
