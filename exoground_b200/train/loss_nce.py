"""Symmetric clip-level InfoNCE of the reference's end-to-end trainer (end2end/main_nce.py:30-48):

    sim = v @ t.T / 0.07 ;  loss = CE(sim, arange) + CE(sim.T, arange) ;  top-1 per text / per video

on the kernels of the alignment path: two similarity GEMMs (sim and sim.T, so both cross-entropies are row-wise),
`exg_xent_diag_fwd/bwd` for the log-softmax with the target on the diagonal (the NUM x NUM probabilities are never
kept), and the gradient through four GEMMs.  Same return dict as the reference's `get_loss`."""
import torch

from .. import ops

INV_TEMP = 1.0 / 0.07


class _SymInfoNceFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, v, t, inv_temp):
        K = ops.get()
        n = v.shape[0]
        s_vt = K.gemm(v, t, alpha=inv_temp, out_dtype=torch.float32)             # [n, n]  rows = videos
        s_tv = K.gemm(t, v, alpha=inv_temp, out_dtype=torch.float32)             # its transpose, rows = texts
        l_v, lse_v = K.xent_diag_fwd(s_vt, n)
        l_t, lse_t = K.xent_diag_fwd(s_tv, n)
        ctx.saved = (v, t, s_vt, s_tv, lse_v, lse_t, inv_temp)
        # F.cross_entropy(sim, target): rows of sim are videos -> the reference calls it loss_per_t (sic)
        return l_v.mean(), l_t.mean(), s_vt

    @staticmethod
    def backward(ctx, g_v, g_t, _):
        K = ops.get()
        v, t, s_vt, s_tv, lse_v, lse_t, inv_temp = ctx.saved
        n = v.shape[0]
        G1 = K.xent_diag_bwd(s_vt, lse_v, inv_temp / n, n, g_dtype=v.dtype)       # d mean CE(sim)   / d (v t^T)
        G2 = K.xent_diag_bwd(s_tv, lse_t, inv_temp / n, n, g_dtype=v.dtype)       # d mean CE(sim^T) / d (t v^T)
        dv = K.gemm(G1, t, transB=True).float() * g_v + K.gemm(G2, t, transA=True, transB=True).float() * g_t
        dt = K.gemm(G1, v, transA=True, transB=True).float() * g_v + K.gemm(G2, v, transB=True).float() * g_t
        return dv.to(v.dtype), dt.to(t.dtype), None


def get_loss(v_features, t_features, device=None, args=None):
    """Drop-in for end2end/main_nce.py::get_loss (device is implied by the tensors)."""
    inv_temp = INV_TEMP if (args is None or getattr(args, "sim", "cos") == "cos") else 1.0
    if not v_features.is_cuda and ops.get().name == "cuda":
        raise ValueError("exoground_b200 runs on CUDA only (no CPU fallback); move inputs to the GPU")
    v, t = v_features.contiguous(), t_features.contiguous()
    loss_a, loss_b, sim = _SymInfoNceFn.apply(v, t, inv_temp)
    target = torch.arange(v.shape[0], device=v.device)
    sim = sim.detach()
    return {'loss': loss_a + loss_b,
            'loss-per-text': loss_a.detach(),
            'loss-per-video': loss_b.detach(),
            'top1-per-text': (sim.argmax(-1) == target).float().mean(),
            'top1-per-video': (sim.argmax(0) == target).float().mean()}
