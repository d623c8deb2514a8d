"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Batch-first, functional, plain-torch fp32 restatement of the reference's
transformer blocks (model/tfm_model.py).  Weights are read from a flat
``state_dict`` using the reference's key names, so a reference checkpoint drives
the oracle directly.  Third-party arithmetic restated here: torch
``nn.MultiheadAttention`` (scaled dot-product attention with an additive -inf
key-padding mask), ``nn.LayerNorm`` (eps 1e-5, biased variance), ``nn.Linear``.
The reference pins no torch version; the oracle is pinned to this image's
torch 2.11.0+cu128 through tests/golden/.
"""
import math
import torch

HEADS = 8  # model/tan_model.py:37-40, model/exo_ground_model.py:56-70 (heads=8 hard-wired)


def layer_norm(x, w, b, eps=1e-5):
    """nn.LayerNorm over the last dim (model/tfm_model.py:22,28)."""
    mu = x.mean(dim=-1, keepdim=True)
    var = ((x - mu) ** 2).mean(dim=-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * w + b


def quick_gelu(x):
    """model/tfm_model.py:11-13."""
    return x * torch.sigmoid(1.702 * x)


def mha(xq, xkv, in_w, in_b, out_w, out_b, key_pad=None, heads=HEADS):
    """nn.MultiheadAttention(q, k=v=xkv, key_padding_mask) restated batch-first.

    xq [B,Sq,C], xkv [B,Sk,C], key_pad [B,Sk] bool (True = ignore key).
    Call sites: model/tfm_model.py:32 (self), :75 (decoder self), :79 (cross).
    """
    B, Sq, C = xq.shape
    Sk = xkv.shape[1]
    d = C // heads
    wq, wk, wv = in_w[:C], in_w[C:2 * C], in_w[2 * C:]
    bq, bk, bv = in_b[:C], in_b[C:2 * C], in_b[2 * C:]
    q = (xq @ wq.T + bq).view(B, Sq, heads, d).transpose(1, 2)     # B,H,Sq,d
    k = (xkv @ wk.T + bk).view(B, Sk, heads, d).transpose(1, 2)
    v = (xkv @ wv.T + bv).view(B, Sk, heads, d).transpose(1, 2)
    s = (q @ k.transpose(-1, -2)) / math.sqrt(d)                    # B,H,Sq,Sk
    if key_pad is not None:
        s = s.masked_fill(key_pad[:, None, None, :].bool(), float("-inf"))
    p = torch.softmax(s, dim=-1)
    o = (p @ v).transpose(1, 2).reshape(B, Sq, C)
    return o @ out_w.T + out_b


def encoder_block(x, sd, pre, key_pad):
    """ResidualAttentionBlock_Step.forward (model/tfm_model.py:34-38).

    Returns (x_out, x_norm) where x_norm = ln_1(x_in)."""
    xn = layer_norm(x, sd[pre + "ln_1.weight"], sd[pre + "ln_1.bias"])
    x = x + mha(xn, xn, sd[pre + "attn.in_proj_weight"], sd[pre + "attn.in_proj_bias"],
                sd[pre + "attn.out_proj.weight"], sd[pre + "attn.out_proj.bias"], key_pad)
    h = layer_norm(x, sd[pre + "ln_2.weight"], sd[pre + "ln_2.bias"])
    h = quick_gelu(h @ sd[pre + "mlp.c_fc.weight"].T + sd[pre + "mlp.c_fc.bias"])
    x = x + h @ sd[pre + "mlp.c_proj.weight"].T + sd[pre + "mlp.c_proj.bias"]
    return x, xn


def temporal_encoder(x, sd, pre, layers, key_pad):
    """TemporalEncoder.forward (model/tfm_model.py:48-55), batch-first.

    Returns the list [ln_1^{(2)}(x_1), ..., ln_1^{(L)}(x_{L-1}), x_L]  (L entries):
    the first block's x_norm is dropped and the raw last output appended."""
    inter = []
    for i in range(layers):
        x, xn = encoder_block(x, sd, f"{pre}resblocks.{i}.", key_pad)
        inter.append(xn)
    inter.pop(0)
    inter.append(x)
    return inter


def decoder_block(x, mem, sd, pre, tgt_pad, mem_pad):
    """ResidualDecoderBlock_Step.forward (model/tfm_model.py:81-86)."""
    xn = layer_norm(x, sd[pre + "ln_1.weight"], sd[pre + "ln_1.bias"])
    x = x + mha(xn, xn, sd[pre + "self_attn.in_proj_weight"], sd[pre + "self_attn.in_proj_bias"],
                sd[pre + "self_attn.out_proj.weight"], sd[pre + "self_attn.out_proj.bias"], tgt_pad)
    h = layer_norm(x, sd[pre + "ln_2.weight"], sd[pre + "ln_2.bias"])
    x = x + mha(h, mem, sd[pre + "attn.in_proj_weight"], sd[pre + "attn.in_proj_bias"],
                sd[pre + "attn.out_proj.weight"], sd[pre + "attn.out_proj.bias"], mem_pad)
    h = layer_norm(x, sd[pre + "ln_3.weight"], sd[pre + "ln_3.bias"])
    h = quick_gelu(h @ sd[pre + "mlp.c_fc.weight"].T + sd[pre + "mlp.c_fc.bias"])
    x = x + h @ sd[pre + "mlp.c_proj.weight"].T + sd[pre + "mlp.c_proj.bias"]
    return x, xn


def temporal_decoder(x, mem, sd, pre, layers, tgt_pad, mem_pad):
    """TemporalDecoder.forward (model/tfm_model.py:96-103), batch-first."""
    inter = []
    for i in range(layers):
        x, xn = decoder_block(x, mem, sd, f"{pre}resblocks.{i}.", tgt_pad, mem_pad)
        inter.append(xn)
    inter.pop(0)
    inter.append(x)
    return inter


def position_embedding_sine(feature_dim=512, num_features=1024, temperature=10000):
    """get_position_embedding_sine (model/tfm_model.py:137-148)."""
    pos = torch.arange(num_features, dtype=torch.float32)
    pos = pos / (pos[-1:] + 1e-6) * (2 * math.pi)
    i = torch.arange(feature_dim, dtype=torch.float32)
    div = temperature ** (2 * torch.div(i, 2, rounding_mode="floor") / feature_dim)
    ang = pos[:, None] / div
    out = torch.stack((ang[:, 0::2].sin(), ang[:, 1::2].cos()), dim=2).flatten(1)
    return out


def interp_pos(table, k, size):
    """F.interpolate(table[None,:k].T, size, 'linear', align_corners=False)
    (model/tan_model.py:151-154) restated as explicit two-tap lerp."""
    src = table[:k]
    scale = k / size
    x = (torch.arange(size, dtype=torch.float32, device=table.device) + 0.5) * scale - 0.5
    x = x.clamp(min=0)
    i0 = x.floor().long().clamp(max=k - 1)
    i1 = (i0 + 1).clamp(max=k - 1)
    w1 = (x - i0.float())[:, None]
    return src[i0] * (1 - w1) + src[i1] * w1
