"""autograd.Function wrappers around single libexg.so ops used outside the transformer stacks:
projections, LayerNorm, positional add / joint concat, L2 normalisation + split, the all-pairs
similarity GEMM (alignment-logit head) and the MIL-NCE reduction.  Each forward/backward is a
handful of `ops.get()` kernel calls; no torch math on activation-sized tensors.
"""
import torch

from .. import ops


def _T(t, T):
    """activation in compute dtype (one cast kernel if needed)."""
    return t if t.dtype == T else ops.get().cast(t.contiguous(), T)


class CastFn(torch.autograd.Function):
    """Differentiable dtype conversion (one cast kernel each way)."""

    @staticmethod
    def forward(ctx, x, dtype):
        ctx.src = x.dtype
        return x if x.dtype == dtype else ops.get().cast(x.contiguous(), dtype)

    @staticmethod
    def backward(ctx, dy):
        return (dy if dy.dtype == ctx.src else ops.get().cast(dy.contiguous(), ctx.src)), None


class LinearFn(torch.autograd.Function):
    """y = x @ W^T (+ b)  (nn.Linear; tan_model.py:42-43,63-66).  x [M,K] in T, W fp32 param with
    compute copy wT (same tensor in fp32 mode), y in out_dtype."""

    @staticmethod
    def forward(ctx, x, w, wT, bias, out_dtype):
        y = ops.get().gemm(x, wT, bias=bias, out_dtype=out_dtype)
        ctx.saved = (x, wT)
        ctx.has_bias = bias is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        K = ops.get()
        x, wT = ctx.saved
        dy = _T(dy.contiguous(), x.dtype)
        dW = K.gemm(dy, x, transA=True, transB=True, out_dtype=torch.float32)
        db = K.colsum(dy) if ctx.has_bias else None
        dx = K.gemm(dy, wT, transB=True) if ctx.needs_input_grad[0] else None
        return dx, dW, None, db, None


class LayerNormFn(torch.autograd.Function):
    """y = LN(x) (* gamma + beta), rows [rows, C]; y in out_dtype; dx returned in x.dtype."""

    @staticmethod
    def forward(ctx, x, g, b, out_dtype):
        y, _ = ops.get().layernorm_fwd(x, g, b, out_dtype=out_dtype)
        ctx.saved = (x, g)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, g = ctx.saved
        K = ops.get()
        want_T = x.dtype != torch.float32
        dx, dxT, dg, db = K.layernorm_bwd(dy.contiguous(), x, g, dxT_dtype=x.dtype if want_T else None,
                                          want_dx=not want_T)
        return (dxT if want_T else dx), dg, db, None


class AddPosFn(torch.autograd.Function):
    """out[b*S+t] = x[b*S+t] + pos[t]  (video_embed + ln_position_init(pos): tan_model.py:161). fp32."""

    @staticmethod
    def forward(ctx, x, pos, B, S):
        out, _ = ops.get().add_rows(x, B * S, add=pos, add_period=S)
        ctx.dims = (B, S)
        return out

    @staticmethod
    def backward(ctx, dy):
        B, S = ctx.dims
        dy = dy.contiguous()
        dpos = ops.get().colsum(dy.view(B, -1)).view(S, -1)
        return dy, dpos, None, None


class JointConcatFn(torch.autograd.Function):
    """x0[b, :Tv] = xv[b] + pos ; x0[b, Tv:] = xt[b]   -> fp32 [B*(Tv+N), C]
    (torch.cat((video_embed_with_time, lang_embed), dim=1): tan_model.py:193-196).  pos may be None."""

    @staticmethod
    def forward(ctx, xv, pos, xt, B, Tv, N):
        K = ops.get()
        Cn = xv.shape[-1]
        out = torch.empty((B * (Tv + N), Cn), dtype=torch.float32, device=xv.device)
        K.add_rows(xv, B * Tv, add=pos, add_period=Tv if pos is not None else 0, out=out, map_out=(Tv, Tv + N, 0))
        K.add_rows(xt, B * N, out=out, map_out=(N, Tv + N, Tv))
        ctx.dims = (B, Tv, N, pos is not None)
        return out

    @staticmethod
    def backward(ctx, dy):
        K = ops.get()
        B, Tv, N, has_pos = ctx.dims
        dy = dy.contiguous()
        dxv, _ = K.add_rows(dy, B * Tv, map_x=(Tv, Tv + N, 0))
        dxt, _ = K.add_rows(dy, B * N, map_x=(N, Tv + N, Tv))
        dpos = K.colsum(dxv.view(B, -1)).view(Tv, -1) if has_pos else None
        return dxv, dpos, dxt, None, None, None


class GatherRowsFn(torch.autograd.Function):
    """Contiguous copy of mapped rows of a [*, C] buffer (slicing the text rows / last stage out of a
    joint stage buffer: tan_model.py:203, exo_ground_model.py:165-166,172).  out dtype = out_dtype."""

    @staticmethod
    def forward(ctx, x, rows, rm, out_dtype):
        K = ops.get()
        xf = x if x.dtype == torch.float32 else K.cast(x, torch.float32)
        out, _ = K.add_rows(xf.view(-1, x.shape[-1]), rows, map_x=rm, out_dtype=out_dtype)
        ctx.meta = (x.shape, x.dtype, rows, rm)
        return out

    @staticmethod
    def backward(ctx, dy):
        K = ops.get()
        shape, dtype, rows, rm = ctx.meta
        dyf = dy.contiguous() if dy.dtype == torch.float32 else K.cast(dy.contiguous(), torch.float32)
        dx = torch.zeros(shape, dtype=dtype, device=dy.device)
        K.add_rows(dyf, rows, out=dx.view(-1, shape[-1]), map_out=rm)
        return dx, None, None, None


class L2NormFn(torch.autograd.Function):
    """y = x / ||x||_2 row-wise, no eps (tan_model.py:110-111); x [..., C] contiguous."""

    @staticmethod
    def forward(ctx, x, out_dtype):
        K = ops.get()
        x2 = x.reshape(-1, x.shape[-1])
        y, inv = K.l2norm_fwd(x2, x2.shape[0], out_dtype=out_dtype)
        ctx.saved = (y, inv, x.shape, x.dtype)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        y, inv, shape, dtype = ctx.saved
        dx = torch.empty(shape, dtype=dtype, device=dy.device)
        ops.get().l2norm_bwd(dy.contiguous().view(-1, shape[-1]), y, inv, dx.view(-1, shape[-1]))
        return dx, None


class SplitL2NormFn(torch.autograd.Function):
    """(v̂ [S,B*Tv,C], t̂ [S,B*N,C]) = L2-normalised video / text rows of joint stages [S, B*(Tv+N), C]
    (tan_model.py:203 + :130-131) in one pass each; backward writes both row sets of one dx buffer."""

    @staticmethod
    def forward(ctx, stages, B, Tv, N):
        K = ops.get()
        S, _, Cn = stages.shape
        flat = stages.view(-1, Cn)
        v, iv = K.l2norm_fwd(flat, S * B * Tv, map_x=(Tv, Tv + N, 0))
        t, it = K.l2norm_fwd(flat, S * B * N, map_x=(N, Tv + N, Tv))
        ctx.saved = (v, iv, t, it, stages.shape, stages.dtype, (B, Tv, N))
        return v.view(S, B * Tv, Cn), t.view(S, B * N, Cn)

    @staticmethod
    def backward(ctx, dv, dt):
        K = ops.get()
        v, iv, t, it, shape, dtype, (B, Tv, N) = ctx.saved
        Cn = shape[-1]
        dx = torch.empty(shape, dtype=dtype, device=dv.device)
        K.l2norm_bwd(dv.contiguous().view(-1, Cn), v, iv, dx.view(-1, Cn), map_dx=(Tv, Tv + N, 0))
        K.l2norm_bwd(dt.contiguous().view(-1, Cn), t, it, dx.view(-1, Cn), map_dx=(N, Tv + N, Tv))
        return dx, None, None, None


class SimLogitsFn(torch.autograd.Function):
    """L[s] = a[s] @ b[s]^T : the all-pairs text-by-time alignment logits
    (einsum "astc,bkc->astbk" / "astc,bskc->astbk": tan_model.py:112-113,132-133).
    a [S,R,C], b [S,Cc,C] or [Cc,C] (shared by all stages, dual branch) -> [S,R,Cc] in a.dtype."""

    @staticmethod
    def forward(ctx, a, b):
        K = ops.get()
        ctx.saved = (a, b)
        return K.gemm(a, b)

    @staticmethod
    def backward(ctx, G):
        K = ops.get()
        a, b = ctx.saved
        G = _T(G.contiguous(), a.dtype)
        S, R, Cn = a.shape
        da = K.gemm(G, b, transB=True)                                  # [S,R,C]
        if b.dim() == 2:   # shared text features: contract over stages and rows at once
            db = K.gemm(G.view(S * R, -1), a.view(S * R, Cn), transA=True, transB=True, out_dtype=torch.float32)
            db = _T(db, b.dtype)
        else:
            db = K.gemm(G, a, transA=True, transB=True)
        return da, db


class NceLossFn(torch.autograd.Function):
    """loss = sum_{s,r} wr[r] (log row_all - log row_pos) + sum_{s,c} wc[c] (log col_all - log col_pos)
    over logits [S,R,Cc]; the exp-sums come from exg_nce_fwd (train/loss.py:251-258), the gradient from
    exg_nce_bwd.  wr / wc already hold the 1/(S*count)/2 factors of the masked means."""

    @staticmethod
    def forward(ctx, logits, sums, tgt, col_valid, wr, wc, dims, inv_temp):
        ra, rp, ca, cp = sums
        tiny = 1e-38
        v = torch.log(ra) - torch.log(rp.clamp(min=tiny))
        t = torch.log(ca.clamp(min=tiny)) - torch.log(cp.clamp(min=tiny))
        loss = (torch.where(wr != 0, v * wr, torch.zeros_like(v))).sum() + \
               (torch.where(wc != 0, t * wc, torch.zeros_like(t))).sum()
        ctx.saved = (logits, sums, tgt, col_valid, wr, wc, dims, inv_temp)
        return loss

    @staticmethod
    def backward(ctx, g):
        logits, sums, tgt, col_valid, wr, wc, (B, T, N), inv_temp = ctx.saved
        G = ops.get().nce_bwd(logits, tgt, col_valid, sums, (wr * g).contiguous(), (wc * g).contiguous(),
                              B, T, N, inv_temp)
        return G, None, None, None, None, None, None, None
