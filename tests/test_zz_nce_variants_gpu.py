"""GPU parity of the remaining NCE variants: the symmetric clip-level InfoNCE of
end2end/main_nce.py:30-48 through the CUDA kernels against the reference formula (two F.cross_entropy calls)."""
import pytest
import torch
import torch.nn.functional as F

from exoground_b200.train import loss_nce
from exoground_b200.utils import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype,n,d", [(torch.float32, 64, 64), (torch.float32, 37, 72), (torch.bfloat16, 256, 512)])
def test_symmetric_infonce(dtype, n, d):
    g = torch.Generator().manual_seed(3)
    v = F.normalize(torch.randn(n, d, generator=g), dim=-1).cuda()
    t = F.normalize(torch.randn(n, d, generator=g), dim=-1).cuda()
    vr, tr = v.clone().requires_grad_(True), t.clone().requires_grad_(True)
    sim = vr.matmul(tr.t()).div(0.07)
    target = torch.arange(n, device="cuda")
    ref = F.cross_entropy(sim, target) + F.cross_entropy(sim.transpose(0, 1), target)
    gv, gt = torch.autograd.grad(ref, (vr, tr))
    v2, t2 = v.to(dtype).requires_grad_(True), t.to(dtype).requires_grad_(True)
    L = loss_nce.get_loss(v2, t2, None, synth.tan_args())
    fp32 = dtype == torch.float32
    tol = 1e-4 if fp32 else 2e-2
    assert abs(float(L["loss"]) - float(ref)) <= tol * max(1.0, abs(float(ref)))
    L["loss"].backward()
    for a, b, name in ((v2.grad, gv, "dv"), (t2.grad, gt, "dt")):
        err = (a.float() - b).norm().item() / b.norm().item()
        assert err <= (1e-3 if fp32 else 5e-2), (name, err)
    if fp32:
        assert float(L["top1-per-text"]) == float((sim.argmax(-1) == target).float().mean())
