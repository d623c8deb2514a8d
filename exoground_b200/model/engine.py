"""Transformer-stack engine: forward and hand-chained backward of the reference's encoder / decoder
stacks (model/tfm_model.py:17-103) expressed as sequences of libexg.so kernels.

Layout: tokens are batch-major rows [B*S, C]; the residual stream is fp32, GEMM operands are in the
compute dtype T (fp32 or bf16).  Stage outputs (the reference's `intermediate` list,
tfm_model.py:49-55) are written by the LayerNorm kernels straight into one [L, B*S, C] buffer, so
there is no torch.stack / permute copy (tan_model.py:170,202).

One torch.autograd.Function per stack keeps Python/autograd overhead at O(1) per stack instead of
O(ops); every arithmetic step is an `ops.get()` call (CUDA kernels; no torch math).
"""
import torch

from .. import ops

C = 512
HEADS = 8
HEAD_DIM = 64
SCALE = HEAD_DIM ** -0.5

ENC_KEYS = ("attn.in_proj_weight", "attn.in_proj_bias", "attn.out_proj.weight", "attn.out_proj.bias",
            "ln_1.weight", "ln_1.bias", "mlp.c_fc.weight", "mlp.c_fc.bias", "mlp.c_proj.weight",
            "mlp.c_proj.bias", "ln_2.weight", "ln_2.bias")
DEC_KEYS = ENC_KEYS + ("self_attn.in_proj_weight", "self_attn.in_proj_bias", "self_attn.out_proj.weight",
                       "self_attn.out_proj.bias", "ln_3.weight", "ln_3.bias")


def _w(p, shadow, T):
    """compute-dtype view of a weight: the fp32 parameter itself, or its bf16 shadow copy."""
    if T == torch.float32:
        return p
    if shadow is None:
        raise RuntimeError("bf16 compute needs shadow weights (CompiledParams.refresh)")
    return shadow


def _mask_u8(mask):
    if mask is None:
        return None
    return mask.to(torch.uint8).contiguous()


# ---------------------------------------------------------------------------------- sub-blocks
def _self_attn_fwd(K, xn, in_w, in_b, out_w, out_b, key_pad, B, S):
    """Returns the residual-branch output (delta, compute dtype); the `x + delta` of tfm_model.py:36 is fused
    into the LayerNorm that follows (exg_add_layernorm_fwd), so the GEMM epilogue stays a pure TMA store."""
    qkv, o, lse = _self_attn_core(K, xn, in_w, in_b, key_pad, B, S)
    delta = K.gemm(o, out_w, bias=out_b)
    return delta, (qkv, o, lse)


def _self_attn_core(K, xn, in_w, in_b, key_pad, B, S):
    """qkv projection + attention; the out-projection is the caller's (fused with the add + LayerNorm that follow)."""
    qkv = K.gemm(xn, in_w, bias=in_b)                                            # [M,3C]
    o, lse = K.attn_fwd(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], key_pad, B, HEADS, S, S, SCALE)
    return qkv, o, lse


def _proj_add_ln(K, a, wT, bias, resid, g, b, T, out=None):
    """(LN(resid + a @ wT^T + bias), resid + a @ wT^T + bias): the out-projection of a sub-block, the residual add of
    tfm_model.py:36-37 and the LayerNorm the next sub-block opens with.  bf16: ONE kernel (exg_gemm_add_ln, the row is
    one TMEM accumulator); fp32 parity path: GEMM + exg_add_layernorm_fwd."""
    if T == torch.bfloat16 and K.gemm_add_ln_ok(a, wT):
        return K.gemm_add_ln(a, wT, bias, resid, g, b, out=out)
    d = K.gemm(a, wT, bias=bias)
    return K.add_layernorm_fwd(d, resid, g, b, out_dtype=T, out=out)


def _sk(sinks, i):
    return sinks[i] if sinks is not None else None


def _self_attn_bwd(K, dxT, saved, xn, in_w, out_w, key_pad, B, S, sk=(None, None, None, None), db_done=False):
    """dxT: grad of the block output of the attention residual branch, in T.  Returns
    (dxn T, dW_in, db_in, dW_out, db_out).  sk = gradient sinks (dW_in, db_in, dW_out, db_out): fp32 views of
    the trainer's flat gradient buffer that the kernels write / accumulate into directly (no autograd add)."""
    qkv, o, lse = saved
    db_out = sk[3] if db_done else K.colsum(dxT, out=sk[3])     # db_done: the LayerNorm backward that made dxT summed it
    dW_out = K.gemm(dxT, o, transA=True, transB=True, out_dtype=torch.float32, out=sk[2])
    do = K.gemm(dxT, out_w, transB=True)
    dqkv = torch.empty_like(qkv)
    K.attn_bwd(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], o, do, lse, key_pad,
               dqkv[:, :C], dqkv[:, C:2 * C], dqkv[:, 2 * C:], B, HEADS, S, S, SCALE)
    db_in = K.colsum(dqkv, out=sk[1])
    dW_in = K.gemm(dqkv, xn, transA=True, transB=True, out_dtype=torch.float32, out=sk[0])
    dxn = K.gemm(dqkv, in_w, transB=True)
    return dxn, dW_in, db_in, dW_out, db_out


def _ffn_fwd(K, h, fc_w, fc_b, proj_w, proj_b):
    """Returns the FFN branch output (delta); `x + delta` (tfm_model.py:37) is fused into the next LayerNorm."""
    f, u = K.gemm(h, fc_w, bias=fc_b, epilogue=ops.EPI_QGELU, want_preact=True)
    delta = K.gemm(f, proj_w, bias=proj_b)
    return delta, (u, f)


def _ffn_bwd(K, dxT, saved, h, fc_w, proj_w, sk=(None, None, None, None), db_done=False):
    """sk = gradient sinks (dW_fc, db_fc, dW_proj, db_proj)."""
    u, f = saved
    db_proj = sk[3] if db_done else K.colsum(dxT, out=sk[3])
    dW_proj = K.gemm(dxT, f, transA=True, transB=True, out_dtype=torch.float32, out=sk[2])
    du = K.gemm(dxT, proj_w, transB=True, epilogue=ops.EPI_QGELU_BWD, aux=u)
    db_fc = K.colsum(du, out=sk[1])
    dW_fc = K.gemm(du, h, transA=True, transB=True, out_dtype=torch.float32, out=sk[0])
    dh = K.gemm(du, fc_w, transB=True)
    return dh, dW_fc, db_fc, dW_proj, db_proj


# ---------------------------------------------------------------------------------- encoder stack
class EncoderStackFn(torch.autograd.Function):
    """stages[L, B*S, C] = TemporalEncoder(x0) with the post-encoder LayerNorm applied to the last
    stage (tfm_model.py:48-55 + tan_model.py:167-168).

    args: x0 fp32 [B*S, C]; key_pad uint8 [B,S] or None; meta = (B, S, L, T, shadows) where shadows
    is a list of per-layer dicts of compute-dtype weight copies (or None in fp32);
    then ln_post gamma/beta and 12 fp32 parameters per layer in ENC_KEYS order."""

    @staticmethod
    def forward(ctx, x0, key_pad, meta, g_post, b_post, *params):
        K = ops.get()
        B, S, L, T, shadows = meta[:5]
        ctx.sinks = meta[5] if len(meta) > 5 else None      # per-parameter gradient sinks (ENC_KEYS order) or None
        ctx.hook = (meta[6], meta[7]) if len(meta) > 7 and meta[6] is not None else None
        M = B * S
        stages = torch.empty((L, M, C), dtype=T, device=x0.device)
        saved = []
        x = x0
        for i in range(L):
            (in_w, in_b, out_w, out_b, g1, b1, fc_w, fc_b, pj_w, pj_b, g2, b2) = params[12 * i:12 * i + 12]
            sh = shadows[i] if shadows is not None else {}
            in_wT, out_wT = _w(in_w, sh.get("in_w"), T), _w(out_w, sh.get("out_w"), T)
            fc_wT, pj_wT = _w(fc_w, sh.get("fc_w"), T), _w(pj_w, sh.get("pj_w"), T)
            if i == 0:
                xn, _ = K.layernorm_fwd(x, g1, b1, out_dtype=T)
            else:   # x_i = x_mid_{i-1} + ffn_{i-1}; stage i-1 = ln_1^{(i)}(x_i)  (tfm_model.py:35-38,49-55)
                xn, x = _proj_add_ln(K, f_prev, pj_prev[0], pj_prev[1], x, g1, b1, T, out=stages[i - 1])
            att = _self_attn_core(K, xn, in_wT, in_b, key_pad, B, S)
            h, x_mid = _proj_add_ln(K, att[1], out_wT, out_b, x, g2, b2, T)
            f_prev, u = K.gemm(h, fc_wT, bias=fc_b, epilogue=ops.EPI_QGELU, want_preact=True)
            pj_prev = (pj_wT, pj_b)
            saved.append((x, xn, att, x_mid, h, (u, f_prev), (in_wT, out_wT, fc_wT, pj_wT)))
            x = x_mid
        _, x = _proj_add_ln(K, f_prev, pj_prev[0], pj_prev[1], x, g_post, b_post, T, out=stages[L - 1])
        ctx.saved = saved
        ctx.x_last = x
        ctx.key_pad = key_pad
        ctx.meta = (B, S, L, T)
        ctx.post = (g_post,)
        ctx.params = params
        return stages

    @staticmethod
    def backward(ctx, d_stages):
        K = ops.get()
        B, S, L, T = ctx.meta
        params = ctx.params
        key_pad = ctx.key_pad
        d_stages = d_stages.contiguous()
        TT = None if T == torch.float32 else T
        grads = [None] * (12 * L)
        sinks = ctx.sinks
        # with gradient sinks the LayerNorm backward that PRODUCES a residual-stream gradient also accumulates its
        # column sums into the bias gradient of the Linear feeding that stream (c_proj / out_proj), so those
        # colsum passes disappear
        fused = sinks is not None
        dx, dxT, dg_post, db_post = K.layernorm_bwd(d_stages[L - 1], ctx.x_last, ctx.post[0], dxT_dtype=TT,
                                                    dxsum_out=sinks[12 * (L - 1) + 9] if fused else None)
        for i in reversed(range(L)):
            (in_w, in_b, out_w, out_b, g1, b1, fc_w, fc_b, pj_w, pj_b, g2, b2) = params[12 * i:12 * i + 12]
            x_in, xn, att, x_mid, h, ffn, (in_wT, out_wT, fc_wT, pj_wT) = ctx.saved[i]
            sk = sinks[12 * i:12 * i + 12] if sinks is not None else [None] * 12
            dh, dW_fc, db_fc, dW_pj, db_pj = _ffn_bwd(K, dxT if TT else dx, ffn, h, fc_wT, pj_wT,
                                                      (sk[6], sk[7], sk[8], sk[9]), db_done=fused)
            dxm, dxmT, dg2, db2 = K.layernorm_bwd(dh, x_mid, g2, dres=dx, dxT_dtype=TT,
                                                  dg_out=(sk[10], sk[11]) if sk[10] is not None else None,
                                                  dxsum_out=sk[3] if fused else None)
            dxn, dW_in, db_in, dW_out, db_out = _self_attn_bwd(K, dxmT if TT else dxm, att, xn, in_wT, out_wT,
                                                               key_pad, B, S, (sk[0], sk[1], sk[2], sk[3]),
                                                               db_done=fused)
            dx, dxT, dg1, db1 = K.layernorm_bwd(dxn, x_in, g1, dy2=d_stages[i - 1] if i > 0 else None,
                                                dres=dxm, dxT_dtype=TT if i > 0 else None,
                                                dg_out=(sk[4], sk[5]) if sk[4] is not None else None,
                                                dxsum_out=sinks[12 * (i - 1) + 9] if fused and i > 0 else None)
            if sinks is None:
                grads[12 * i:12 * i + 12] = [dW_in, db_in, dW_out, db_out, dg1, db1, dW_fc, db_fc, dW_pj, db_pj,
                                             dg2, db2]
            ctx.saved[i] = None
            if ctx.hook is not None:        # block i's weight gradients are final in the trainer's flat buffer
                ctx.hook[0](f"{ctx.hook[1]}resblocks.{i}.")
        return (dx, None, None, dg_post, db_post, *grads)


# ---------------------------------------------------------------------------------- decoder stack
class DecoderStackFn(torch.autograd.Function):
    """out[B*N, C] (compute dtype) = last entry of TemporalDecoder(x, memory) (tfm_model.py:89-103):
    the raw output of the last block (exo_ground_model.py:172-173 uses decoder_output[-1] only).

    args: x0 fp32 [B*N, C]; mem T [B*Sm, C]; tgt_pad / mem_pad uint8 or None;
    meta = (B, N, Sm, L, T, shadows); then 18 fp32 parameters per layer in DEC_KEYS order."""

    @staticmethod
    def forward(ctx, x0, mem, tgt_pad, mem_pad, meta, *params):
        K = ops.get()
        B, N, Sm, L, T, shadows = meta
        saved = []
        x = x0
        for i in range(L):
            (cin_w, cin_b, cout_w, cout_b, g1, b1, fc_w, fc_b, pj_w, pj_b, g2, b2,
             sin_w, sin_b, sout_w, sout_b, g3, b3) = params[18 * i:18 * i + 18]
            sh = shadows[i] if shadows is not None else {}
            cin_wT, cout_wT = _w(cin_w, sh.get("in_w"), T), _w(cout_w, sh.get("out_w"), T)
            fc_wT, pj_wT = _w(fc_w, sh.get("fc_w"), T), _w(pj_w, sh.get("pj_w"), T)
            sin_wT, sout_wT = _w(sin_w, sh.get("sin_w"), T), _w(sout_w, sh.get("sout_w"), T)
            if i == 0:
                xn, _ = K.layernorm_fwd(x, g1, b1, out_dtype=T)
            else:
                xn, x = K.add_layernorm_fwd(d3, x, g1, b1, out_dtype=T)
            att = _self_attn_core(K, xn, sin_wT, sin_b, tgt_pad, B, N)
            h2, x1 = _proj_add_ln(K, att[1], sout_wT, sout_b, x, g2, b2, T)
            q = K.gemm(h2, cin_wT[:C], bias=cin_b[:C])                            # [B*N, C]
            kv = K.gemm(mem, cin_wT[C:], bias=cin_b[C:])                          # [B*Sm, 2C]
            o, lse = K.attn_fwd(q, kv[:, :C], kv[:, C:], mem_pad, B, HEADS, N, Sm, SCALE)
            h3, x2 = _proj_add_ln(K, o, cout_wT, cout_b, x1, g3, b3, T)
            d3, ffn = _ffn_fwd(K, h3, fc_wT, fc_b, pj_wT, pj_b)
            saved.append((x, xn, att, x1, h2, q, kv, o, lse, x2, h3, ffn,
                          (cin_wT, cout_wT, fc_wT, pj_wT, sin_wT, sout_wT)))
            x = x2
        ctx.saved, ctx.mem, ctx.pads, ctx.meta, ctx.params = saved, mem, (tgt_pad, mem_pad), (B, N, Sm, L, T), params
        # raw output of the last block: x2 + ffn (tfm_model.py:86), in the compute dtype
        d3f = d3 if d3.dtype == torch.float32 else K.cast(d3, torch.float32)
        out, _ = K.add_rows(x, B * N, x2=d3f, out_dtype=T)
        return out

    @staticmethod
    def backward(ctx, d_out):
        K = ops.get()
        B, N, Sm, L, T = ctx.meta
        params, mem = ctx.params, ctx.mem
        tgt_pad, mem_pad = ctx.pads
        TT = None if T == torch.float32 else T
        grads = [None] * (18 * L)
        d_out = d_out.contiguous()
        dxT = d_out
        dx = K.cast(d_out, torch.float32) if TT else d_out
        dmem = torch.zeros(mem.shape, dtype=torch.float32, device=mem.device)
        for i in reversed(range(L)):
            (cin_w, cin_b, cout_w, cout_b, g1, b1, fc_w, fc_b, pj_w, pj_b, g2, b2,
             sin_w, sin_b, sout_w, sout_b, g3, b3) = params[18 * i:18 * i + 18]
            (x_in, xn, att, x1, h2, q, kv, o, lse, x2, h3, ffn,
             (cin_wT, cout_wT, fc_wT, pj_wT, sin_wT, sout_wT)) = ctx.saved[i]
            dh3, dW_fc, db_fc, dW_pj, db_pj = _ffn_bwd(K, dxT if TT else dx, ffn, h3, fc_wT, pj_wT)
            dx2, dx2T, dg3, db3 = K.layernorm_bwd(dh3, x2, g3, dres=dx, dxT_dtype=TT)
            g2T = dx2T if TT else dx2
            # cross attention
            db_cout = K.colsum(g2T)
            dW_cout = K.gemm(g2T, o, transA=True, transB=True, out_dtype=torch.float32)
            do = K.gemm(g2T, cout_wT, transB=True)
            dq, dkv = torch.empty_like(q), torch.empty_like(kv)
            K.attn_bwd(q, kv[:, :C], kv[:, C:], o, do, lse, mem_pad, dq, dkv[:, :C], dkv[:, C:], B, HEADS, N, Sm, SCALE)
            db_cin = torch.cat((K.colsum(dq), K.colsum(dkv)))
            dW_cin = torch.empty((3 * C, C), dtype=torch.float32, device=q.device)
            K.gemm(dq, h2, transA=True, transB=True, out_dtype=torch.float32, out=dW_cin[:C])
            K.gemm(dkv, mem, transA=True, transB=True, out_dtype=torch.float32, out=dW_cin[C:])
            dh2 = K.gemm(dq, cin_wT[:C], transB=True)
            K.gemm(dkv, cin_wT[C:], transB=True, epilogue=ops.EPI_RESID, aux=dmem, out=dmem, out_dtype=torch.float32)
            dx1, dx1T, dg2, db2 = K.layernorm_bwd(dh2, x1, g2, dres=dx2, dxT_dtype=TT)
            dxn, dW_sin, db_sin, dW_sout, db_sout = _self_attn_bwd(K, dx1T if TT else dx1, att, xn, sin_wT,
                                                                   sout_wT, tgt_pad, B, N)
            dx, dxT, dg1, db1 = K.layernorm_bwd(dxn, x_in, g1, dres=dx1, dxT_dtype=TT if i > 0 else None)
            grads[18 * i:18 * i + 18] = [dW_cin, db_cin, dW_cout, db_cout, dg1, db1, dW_fc, db_fc, dW_pj, db_pj,
                                         dg2, db2, dW_sin, db_sin, dW_sout, db_sout, dg3, db3]
            ctx.saved[i] = None
        dmemT = K.cast(dmem, T) if TT else dmem
        return (dx, dmemT, None, None, None, *grads)
