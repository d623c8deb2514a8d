"""Drop-in mirror of the reference's model/tfm_model.py: same module tree and parameter names
(`resblocks.{i}.attn.in_proj_weight`, `.ln_1.weight`, `.mlp.c_fc.weight`, ...), so reference
checkpoints load unchanged.  The torch modules below are *parameter holders* only: every forward goes
through exoground_b200.model.engine (CUDA kernels), never through nn.MultiheadAttention / nn.Linear math.
"""
import math
from collections import OrderedDict

import torch
from torch import nn
from torch.nn import LayerNorm

from . import engine


class QuickGELU(nn.Module):
    """x * sigmoid(1.702 x) (reference tfm_model.py:11-13); fused into the c_fc GEMM epilogue."""

    def forward(self, x):
        raise RuntimeError("QuickGELU is fused into the c_fc GEMM epilogue (EXG_EPI_QGELU)")


class ResidualAttentionBlock_Step(nn.Module):
    """Parameter holder for one encoder block (reference tfm_model.py:17-38)."""

    def __init__(self, d_model, n_head):
        super().__init__()
        self.attn = nn.MultiheadAttention(d_model, n_head)
        self.ln_1 = LayerNorm(d_model)
        self.mlp = nn.Sequential(OrderedDict([("c_fc", nn.Linear(d_model, d_model * 4)), ("gelu", QuickGELU()),
                                              ("c_proj", nn.Linear(d_model * 4, d_model))]))
        self.ln_2 = LayerNorm(d_model)


class ResidualDecoderBlock_Step(nn.Module):
    """Parameter holder for one decoder block (reference tfm_model.py:59-86)."""

    def __init__(self, d_model, n_head):
        super().__init__()
        self.self_attn = nn.MultiheadAttention(d_model, n_head)
        self.ln_1 = LayerNorm(d_model)
        self.attn = nn.MultiheadAttention(d_model, n_head)
        self.mlp = nn.Sequential(OrderedDict([("c_fc", nn.Linear(d_model, d_model * 4)), ("gelu", QuickGELU()),
                                              ("c_proj", nn.Linear(d_model * 4, d_model))]))
        self.ln_2 = LayerNorm(d_model)
        self.ln_3 = LayerNorm(d_model)


_SHADOW = {"attn.in_proj_weight": "in_w", "attn.out_proj.weight": "out_w", "mlp.c_fc.weight": "fc_w",
           "mlp.c_proj.weight": "pj_w", "self_attn.in_proj_weight": "sin_w", "self_attn.out_proj.weight": "sout_w"}


def _stack_args(module, prefix, keys, wT, T):
    """(flat param list in `keys` order per layer, per-layer shadow dicts or None)."""
    named = dict(module.named_parameters())
    params, shadows = [], ([] if T != torch.float32 else None)
    for i in range(module.layers):
        for k in keys:
            params.append(named[f"resblocks.{i}.{k}"])
        if shadows is not None:
            shadows.append({short: wT(f"{prefix}resblocks.{i}.{k}") for k, short in _SHADOW.items() if k in keys})
    return params, shadows


class TemporalEncoder(nn.Module):
    def __init__(self, width, layers, heads):
        super().__init__()
        if width != engine.C or heads != engine.HEADS:
            raise ValueError("the sm_100a kernels are specialised for width 512 / 8 heads "
                             "(reference tan_model.py:37-40, exo_ground_model.py:56-70)")
        self.width, self.layers = width, layers
        self.resblocks = nn.ModuleList([ResidualAttentionBlock_Step(width, heads) for _ in range(layers)])

    def run(self, x0, key_pad_u8, B, S, wT, prefix, g_post, b_post, T):
        """x0 fp32 [B*S, C] -> stages [L, B*S, C] (T); last stage has (g_post, b_post) LayerNorm applied."""
        params, shadows = _stack_args(self, prefix, engine.ENC_KEYS, wT, T)
        sinks = None
        if getattr(self, "_grad_sink", False) and torch.is_grad_enabled():
            # a trainer owns (and zeroes) the .grad views: the backward kernels write into them directly
            if all(p.grad is not None and p.grad.dtype == torch.float32 and p.grad.is_contiguous() for p in params):
                sinks = [p.grad for p in params]
        # a data-parallel trainer may ask to be told when a block's gradients are final (overlapped all-reduce)
        hook = getattr(self, "_grad_ready_hook", None) if sinks is not None else None
        return engine.EncoderStackFn.apply(x0, key_pad_u8, (B, S, self.layers, T, shadows, sinks, hook, prefix),
                                           g_post, b_post, *params)


class TemporalDecoder(nn.Module):
    def __init__(self, width, layers, heads):
        super().__init__()
        if width != engine.C or heads != engine.HEADS:
            raise ValueError("the sm_100a kernels are specialised for width 512 / 8 heads")
        self.width, self.layers = width, layers
        self.resblocks = nn.ModuleList([ResidualDecoderBlock_Step(width, heads) for _ in range(layers)])

    def run(self, x0, mem, tgt_pad_u8, mem_pad_u8, B, N, Sm, wT, prefix, T):
        """x0 fp32 [B*N, C], mem T [B*Sm, C] -> last block output [B*N, C] (T)."""
        params, shadows = _stack_args(self, prefix, engine.DEC_KEYS, wT, T)
        return engine.DecoderStackFn.apply(x0, mem, tgt_pad_u8, mem_pad_u8, (B, N, Sm, self.layers, T, shadows), *params)


def get_position_embedding_sine(feature_dim=512, num_features=1024, temperature=10000):
    """Sine table [num_features, feature_dim] (reference tfm_model.py:137-148); host-side, build time."""
    pos = torch.arange(num_features, dtype=torch.float32)
    pos = pos / (pos[-1:] + 1e-6) * (2 * math.pi)
    i = torch.arange(feature_dim, dtype=torch.float32)
    div = temperature ** (2 * torch.div(i, 2, rounding_mode="floor") / feature_dim)
    ang = pos[:, None] / div
    return torch.stack((ang[:, 0::2].sin(), ang[:, 1::2].cos()), dim=2).flatten(1)


def interp_rows(table, k, size):
    """Linear interpolation of the first k rows of a pos-embed table to `size` rows, identical to
    F.interpolate(mode='linear', align_corners=False) (reference tan_model.py:151-154).  Runs on a
    [<=4096, 512] table once per forward: host-side plumbing, differentiable w.r.t. the table."""
    scale = k / size
    x = ((torch.arange(size, dtype=torch.float32, device=table.device) + 0.5) * scale - 0.5).clamp(min=0)
    i0 = x.floor().long().clamp(max=k - 1)
    i1 = (i0 + 1).clamp(max=k - 1)
    w1 = (x - i0.float())[:, None]
    src = table[:k]
    return src[i0] * (1 - w1) + src[i1] * w1
