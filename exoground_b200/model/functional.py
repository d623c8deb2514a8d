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


class RowDotFn(torch.autograd.Function):
    """y [M] = x @ w[0] + b for a one-output nn.Linear (the alignability head, tan_model.py:64-68,128-134):
    a GEMM with N = 1 is pure HBM traffic, so forward, dx and dw are row-wise kernels.  x [M,C] in T; fp32 y."""

    @staticmethod
    def forward(ctx, x, w, bias):
        ctx.saved = (x, w)
        ctx.has_bias = bias is not None
        return ops.get().rowdot(x, w.reshape(-1), bias)

    @staticmethod
    def backward(ctx, dy):
        K = ops.get()
        x, w = ctx.saved
        dy = dy.contiguous().float()
        dw = K.wcolsum(x, dy).view(1, -1)
        db = dy.sum().reshape(1) if ctx.has_bias else None
        dx = K.outer_rows(dy, w.reshape(-1), x.dtype) if ctx.needs_input_grad[0] else None
        return dx, dw, db


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


class JointConcat3Fn(torch.autograd.Function):
    """x0[b] = [xv[b] | xa[b] + pos | xt[b]]  -> fp32 [B*(2Tv+N), C]
    (torch.cat((video, audio + ln_position_init(pos), lang), dim=1): exo_ground_model.py:347-350)."""

    @staticmethod
    def forward(ctx, xv, xa, pos, xt, B, Tv, N):
        K = ops.get()
        S = 2 * Tv + N
        out = torch.empty((B * S, xv.shape[-1]), dtype=torch.float32, device=xv.device)
        K.add_rows(xv, B * Tv, out=out, map_out=(Tv, S, 0))
        K.add_rows(xa, B * Tv, add=pos, add_period=Tv, out=out, map_out=(Tv, S, Tv))
        K.add_rows(xt, B * N, out=out, map_out=(N, S, 2 * Tv))
        ctx.dims = (B, Tv, N)
        return out

    @staticmethod
    def backward(ctx, dy):
        K = ops.get()
        B, Tv, N = ctx.dims
        S = 2 * Tv + N
        dy = dy.contiguous()
        dxv, _ = K.add_rows(dy, B * Tv, map_x=(Tv, S, 0))
        dxa, _ = K.add_rows(dy, B * Tv, map_x=(Tv, S, Tv))
        dxt, _ = K.add_rows(dy, B * N, map_x=(N, S, 2 * Tv))
        dpos = K.colsum(dxa.view(B, -1)).view(Tv, -1)
        return dxv, dxa, dpos, dxt, None, None, None


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


def nce_value(sums, wr, wc, col_valid=None, inv_temp=0.0, masked=False, n_rows=None):
    """sum_r wr (log row_all - log row_pos) + sum_c wc (log col_all - log col_pos) from the exp-sums.
    masked=True (self-labelling branch): a weighted row / column whose positives were ALL removed by the own-row
    mask gets what the reference computes there — its positive LSE is over -6e4 fills only, i.e.
    -6e4 + log(#entries) (train/loss.py:100-106 + :247-256), so the term is LSE_all + 6e4 - log(#entries)."""
    ra, rp, ca, cp = sums
    tiny = 1e-38
    v = torch.log(ra) - torch.log(rp.clamp(min=tiny))
    t = torch.log(ca.clamp(min=tiny)) - torch.log(cp.clamp(min=tiny))
    if masked:
        n_cols = col_valid.float().sum()
        n_rows = float(ra.shape[-1] if n_rows is None else n_rows)    # rows of the GLOBAL logits
        v = torch.where(rp > 0, v, torch.log(ra) + inv_temp + 6e4 - torch.log(n_cols))
        t = torch.where(cp > 0, t, torch.log(ca.clamp(min=tiny)) + inv_temp + 6e4 - torch.log(torch.tensor(n_rows)))
    z = torch.zeros((), device=ra.device)
    return torch.where(wr != 0, v * wr, z).sum() + torch.where(wc != 0, t * wc, z).sum()


class NceLossFn(torch.autograd.Function):
    """loss = sum_{s,r} wr[r] (log row_all - log row_pos) + sum_{s,c} wc[c] (log col_all - log col_pos)
    over logits [S,R,Cc]; the exp-sums come from exg_nce_fwd (train/loss.py:251-258), the gradient from
    exg_nce_bwd.  wr / wc already hold the 1/(S*count)/2 factors of the masked means."""

    @staticmethod
    def forward(ctx, logits, sums, tgt, col_valid, wr, wc, dims, inv_temp, own_cols=None):
        # dims = (B, T, N) or (B, T, N, Bc, clip0).  own_cols (cross-GPU negatives, train/xneg.py): 0/1 over the
        # global columns; the VALUE counts only this rank's columns (shares add up to the global loss over ranks),
        # the GRADIENT is that of the global loss with respect to this rank's row block (all columns).
        T_, N_ = dims[1], dims[2]
        Bc = dims[3] if len(dims) > 3 else dims[0]
        loss = nce_value(sums, wr, wc if own_cols is None else wc * own_cols, col_valid, inv_temp,
                         masked=isinstance(tgt, tuple) and len(tgt) > 2 and tgt[2] is not None, n_rows=Bc * T_)
        ctx.saved = (logits, sums, tgt, col_valid, wr, wc, dims, inv_temp)
        return loss

    @staticmethod
    def backward(ctx, g):
        logits, sums, tgt, col_valid, wr, wc, dims, inv_temp = ctx.saved
        B, T, N = dims[:3]
        Bc, clip0 = (dims[3], dims[4]) if len(dims) > 3 else (B, 0)
        # (uint8 mask, nce_pack_targets(mask), optional own-row mask of the self-labelling branch)
        tgt, bits, orm = (tuple(tgt) + (None, None))[:3] if isinstance(tgt, tuple) else (tgt, None, None)
        G = ops.get().nce_bwd(logits, tgt, col_valid, sums, (wr * g).contiguous(), (wc * g).contiguous(),
                              B, T, N, inv_temp, tgt_bits=bits, own_row_mask=orm, Bc=Bc, clip0=clip0)
        return G, None, None, None, None, None, None, None, None


class StageMeanFn(torch.autograd.Function):
    """mean over the stage axis of stages [S, M, C] -> fp32 [M, C]
    (`get_unimodal_features(...).mean(dim=1)`: exo_ground_model.py:153-154)."""

    @staticmethod
    def forward(ctx, stages):
        out, _ = ops.get().stage_mean(stages.contiguous())
        ctx.meta = (stages.shape[0], stages.dtype)
        return out

    @staticmethod
    def backward(ctx, g):
        S, dtype = ctx.meta
        return ops.get().bcast_rows(g.contiguous().float(), S, 1.0 / S, dtype)


class InfoNceBmmFn(torch.autograd.Function):
    """-mean_b,t log_softmax(f1n[b] f2n[b]^T / temp)[t, t] for L2-normalised f1n, f2n [B,T,D]
    (compute_info_nce_loss: exo_ground_model.py:206-234, loss_egoexo4d.py:325-347).  One batched GEMM for the
    T x T similarities, one row-wise diagonal cross-entropy kernel; the gradient goes back through two GEMMs."""

    @staticmethod
    def forward(ctx, f1n, f2n, inv_temp):
        K = ops.get()
        B, T, D = f1n.shape
        sim = K.gemm(f1n, f2n, alpha=inv_temp, out_dtype=torch.float32)           # [B,T,T]
        loss_rows, lse = K.xent_diag_fwd(sim.view(B * T, T), T)
        ctx.saved = (f1n, f2n, sim, lse, inv_temp)
        return loss_rows.mean()

    @staticmethod
    def backward(ctx, g):
        K = ops.get()
        f1n, f2n, sim, lse, inv_temp = ctx.saved
        B, T, D = f1n.shape
        # d sim (already holds the 1/(B*T) of the mean; the upstream scalar is applied on the small outputs)
        G = K.xent_diag_bwd(sim.view(B * T, T), lse, inv_temp / (B * T), T, g_dtype=f1n.dtype).view(B, T, T)
        d1 = K.gemm(G, f2n, transB=True) if ctx.needs_input_grad[0] else None      # [B,T,D] = G @ f2n
        d2 = K.gemm(G, f1n, transA=True, transB=True) if ctx.needs_input_grad[1] else None   # G^T @ f1n
        return (d1 * g if d1 is not None else None), (d2 * g if d2 is not None else None), None


class XviewNceFn(torch.autograd.Function):
    """Per-(b,t) cross-view InfoNCE + monitors (loss_egoexo4d.py:269-323, :201-233); differentiable in out_feat."""

    @staticmethod
    def forward(ctx, out_feat, ego, pos_idx, neg_idx, same_feat, same_idx, only_same, inv_temp):
        K = ops.get()
        loss, l1, pc, nc, stats = K.xview_nce_fwd(out_feat.contiguous(), ego, pos_idx, neg_idx, same_feat, same_idx,
                                                  only_same, inv_temp)
        ctx.saved = (out_feat, ego, pos_idx, neg_idx, same_feat, same_idx, only_same, inv_temp, stats)
        ctx.mark_non_differentiable(l1, pc, nc)
        return loss, l1, pc, nc

    @staticmethod
    def backward(ctx, gl, *_):
        out_feat, ego, pos_idx, neg_idx, same_feat, same_idx, only_same, inv_temp, stats = ctx.saved
        go = ops.get().xview_nce_bwd(out_feat.contiguous(), ego, pos_idx, neg_idx, same_feat, same_idx, only_same,
                                     inv_temp, stats, gl.contiguous().float())
        return go, None, None, None, None, None, None, None


class GroundLossFn(torch.autograd.Function):
    """(1 - mean IoU) + L1 terms of get_grounding_loss_reg_head (loss_egoexo4d.py:73-145) in one launch.
    Returns (loss, stats, iou); only `loss` is differentiable (w.r.t. preds [M,2])."""

    @staticmethod
    def forward(ctx, preds, gt_a, gt_b, text_pad_u8, use_cd, eps, thresholds):
        stats, iou, dp = ops.get().ground_loss(preds.contiguous().float(), gt_a, gt_b, text_pad_u8, use_cd, eps,
                                               thresholds)
        l1 = stats[0] + stats[1] if use_cd else (stats[0] + stats[1]) * 0.5
        loss = (1.0 - stats[2]) + l1
        ctx.saved = (dp, preds.dtype)
        ctx.mark_non_differentiable(stats, iou)
        return loss, stats, iou

    @staticmethod
    def backward(ctx, g, *_):
        dp, dtype = ctx.saved
        return (dp * g).to(dtype), None, None, None, None, None, None


class MlpReluFn(torch.autograd.Function):
    """y = relu(x W1^T + b1) W2^T + b2 (ViewInvariantMLP.mlp: vi_encoder.py:215-219).  ReLU lives in the first GEMM's
    epilogue, its mask in the epilogue of the GEMM that produces dh.  x [M,D] in T; y in out_dtype."""

    @staticmethod
    def forward(ctx, x, w1, w1T, b1, w2, w2T, b2, out_dtype):
        from ..ops import lib as L
        K = ops.get()
        h = K.gemm(x, w1T, bias=b1, epilogue=L.EPI_RELU)
        y = K.gemm(h, w2T, bias=b2, out_dtype=out_dtype)
        ctx.saved = (x, h, w1T, w2T)
        return y

    @staticmethod
    def backward(ctx, dy):
        from ..ops import lib as L
        K = ops.get()
        x, h, w1T, w2T = ctx.saved
        dy = _T(dy.contiguous(), x.dtype)
        dW2 = K.gemm(dy, h, transA=True, transB=True, out_dtype=torch.float32)
        db2 = K.colsum(dy)
        dh = K.gemm(dy, w2T, transB=True, epilogue=L.EPI_RELU_BWD, aux=h)
        dW1 = K.gemm(dh, x, transA=True, transB=True, out_dtype=torch.float32)
        db1 = K.colsum(dh)
        dx = K.gemm(dh, w1T, transB=True) if ctx.needs_input_grad[0] else None
        return dx, dW1, None, db1, dW2, None, db2, None
