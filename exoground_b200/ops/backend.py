"""Tensor-level op API on top of the C ABI (exoground_b200/ops/lib.py).

Every method takes/returns torch CUDA tensors, extracts raw pointers / strides and enqueues the
CUDA kernels of libexg.so on torch's current stream.  This is the only place tensors meet ctypes.
The op *specifications* (shapes, dtypes, semantics) are documented here once; tests/fake_backend.py
implements the same spec with plain torch ops on CPU to validate the host logic without a GPU.

Conventions: "T" = compute dtype (torch.float32 or torch.bfloat16); rm = row map tuple
(group, group_stride, offset) or None for contiguous rows (see exg_rowmap in include/exg.h).
"""
import ctypes as C
import os

import torch

from . import lib as L

_DT = {torch.float32: L.F32, torch.bfloat16: L.BF16}


def _dt(t):
    try:
        return _DT[t]
    except KeyError:
        raise ValueError(f"exoground_b200: unsupported dtype {t} (float32 / bfloat16 only)")


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _rm(rm, rows):
    if rm is None:
        return L.RowMap(max(int(rows), 1), max(int(rows), 1), 0)
    return L.RowMap(int(rm[0]), int(rm[1]), int(rm[2]))


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise ValueError("exoground_b200 ops need CUDA tensors (there is no CPU fallback)")


def _rows2d(t):
    """(data tensor, leading dim) of a 2-D operand whose last dim is contiguous."""
    if t.dim() != 2 or (t.shape[1] > 1 and t.stride(1) != 1):
        raise ValueError(f"expected a 2-D operand with contiguous last dim, got {tuple(t.shape)} / {t.stride()}")
    return t.stride(0) if t.shape[0] > 1 else max(t.shape[1], 1)


class CudaBackend:
    name = "cuda"

    def __init__(self):
        self.lib = L.load()

    # ------------------------------------------------------------------ GEMM
    def gemm(self, A, B, *, transA=False, transB=False, bias=None, epilogue=L.EPI_NONE, aux=None,
             out_dtype=None, alpha=1.0, out=None, want_preact=False, impl=L.GEMM_AUTO):
        """C = epi(alpha * op(A) @ op(B) (+ bias)).

        A: [M,K] (transA=False) or [K,M] (transA=True); B: [N,K] (transB=False, nn.Linear layout) or
        [K,N] (transB=True).  Either may carry a leading batch dim; a 2-D operand is broadcast over the
        batch.  bias fp32 [N].  aux: RESID -> fp32 [.., M, N]; *_BWD -> T [.., M, N].
        Returns C [.., M, N] of out_dtype (default: input dtype), plus the pre-activation if want_preact."""
        _cuda(A, B, bias, aux, out)
        if A.dtype != B.dtype:
            raise ValueError("gemm: A and B dtypes differ")
        batch = A.shape[0] if A.dim() == 3 else (B.shape[0] if B.dim() == 3 else 1)
        a2 = A[0] if A.dim() == 3 else A
        b2 = B[0] if B.dim() == 3 else B
        M, K = (a2.shape[1], a2.shape[0]) if transA else (a2.shape[0], a2.shape[1])
        N, Kb = (b2.shape[1], b2.shape[0]) if transB else (b2.shape[0], b2.shape[1])
        if K != Kb:
            raise ValueError(f"gemm: inner dims differ ({K} vs {Kb})")
        out_dtype = out_dtype or A.dtype
        batched = A.dim() == 3 or B.dim() == 3
        if out is None:
            out = torch.empty((batch, M, N) if batched else (M, N), dtype=out_dtype, device=A.device)
        c2 = out[0] if out.dim() == 3 else out
        pre = torch.empty_like(out) if want_preact else None
        p = L.GemmParams()
        p.A, p.B, p.C = A.data_ptr(), B.data_ptr(), out.data_ptr()
        p.M, p.N, p.K = M, N, K
        p.lda, p.ldb, p.ldc = _rows2d(a2), _rows2d(b2), _rows2d(c2)
        p.transA, p.transB = int(transA), int(transB)
        p.batch = batch
        p.strideA = A.stride(0) if A.dim() == 3 else 0
        p.strideB = B.stride(0) if B.dim() == 3 else 0
        p.strideC = out.stride(0) if out.dim() == 3 else 0
        p.in_dtype, p.out_dtype, p.epilogue = _dt(A.dtype), _dt(out.dtype), epilogue
        p.bias = bias.data_ptr() if bias is not None else None
        if aux is not None:
            x2 = aux[0] if aux.dim() == 3 else aux
            p.aux, p.ldaux = aux.data_ptr(), _rows2d(x2)
            p.strideAux = aux.stride(0) if aux.dim() == 3 else 0
        p.aux_out = pre.data_ptr() if pre is not None else None
        p.alpha, p.impl = float(alpha), impl
        L.check(self.lib.exg_gemm(C.byref(p), _stream()), "exg_gemm")
        return (out, pre) if want_preact else out

    def gemm_add_ln_ok(self, A, W):
        """True if (A [M,K] bf16, W [512,K] bf16) should take the fused out-projection + residual add + LayerNorm kernel.
        Opt-in (EXG_GEMM_LN=1): inside the train step it measured SLOWER than GEMM + exg_add_layernorm_fwd (7.86 vs
        7.40 ms per step, profiles/gemm_ln_r02.txt) - the 128 x 512 tile fills TMEM, so MMA and epilogue of a tile cannot
        overlap, only M/128 <= 96 CTAs run, and its thread = row global accesses are uncoalesced."""
        if os.environ.get("EXG_GEMM_LN", "0") != "1" or A.dtype != torch.bfloat16 or W.dtype != torch.bfloat16:
            return False
        if A.dim() != 2 or W.dim() != 2 or W.shape[0] != 512 or A.shape[1] != W.shape[1]:
            return False
        if A.stride(1) != 1 or W.stride(1) != 1:
            return False
        return bool(self.lib.exg_gemm_add_ln_supported(_p(A), _p(W), A.shape[0], A.shape[1], _rows2d(A), _rows2d(W)))

    def gemm_add_ln(self, A, W, bias, resid, gamma, beta, *, out=None, eps=1e-5):
        """s = resid + A @ W^T + bias (fp32, new tensor);  y = LN(s) * gamma + beta (bf16; `out` if given).
        Returns (y, s) like add_layernorm_fwd."""
        _cuda(A, W, bias, resid, gamma, beta, out)
        M, Kd = A.shape
        if out is None:
            out = torch.empty((M, 512), dtype=torch.bfloat16, device=A.device)
        if not (resid.is_contiguous() and out.is_contiguous() and resid.dtype == torch.float32
                and out.dtype == torch.bfloat16 and tuple(resid.shape) == (M, 512) and tuple(out.shape) == (M, 512)):
            raise ValueError("gemm_add_ln: resid fp32 [M,512] / out bf16 [M,512] contiguous expected")
        s = torch.empty((M, 512), dtype=torch.float32, device=A.device)
        L.check(self.lib.exg_gemm_add_ln(_p(A), _p(W), _p(bias), _p(resid), _p(gamma), _p(beta), float(eps), _p(s),
                                         _p(out), M, Kd, _rows2d(A), _rows2d(W), _stream()), "exg_gemm_add_ln")
        return out, s

    # ------------------------------------------------------------------ row-wise
    def layernorm_fwd(self, x, gamma, beta, *, out_dtype, out=None, out_map=None, y2_dtype=None, add=None,
                      add_period=0, eps=1e-5):
        """y[map(r)] = LN(x[r]) (+ add[r % add_period]); x [rows,C] contiguous.  `out` may be a larger
        buffer addressed through out_map.  Returns (y, y2) with y2 an optional contiguous second copy."""
        _cuda(x, gamma, beta, out, add)
        rows, Cn = x.shape
        if out is None:
            out = torch.empty((rows, Cn), dtype=out_dtype, device=x.device)
        y2 = torch.empty((rows, Cn), dtype=y2_dtype, device=x.device) if y2_dtype is not None else None
        L.check(self.lib.exg_layernorm_fwd(_p(x), _dt(x.dtype), _p(gamma), _p(beta), _p(out), _dt(out.dtype),
                                           _rm(out_map, rows), _p(y2), _dt(y2.dtype) if y2 is not None else 0,
                                           _p(add), int(add_period), None, None, rows, Cn, float(eps),
                                           _stream()), "exg_layernorm_fwd")
        return out, y2

    def add_layernorm_fwd(self, x, resid, gamma, beta, *, out_dtype, out=None, out_map=None, eps=1e-5):
        """s = resid + x (fp32 residual stream, new tensor); y[map(r)] = LN(s).  Returns (y, s)."""
        _cuda(x, resid, gamma, beta, out)
        rows, Cn = x.shape
        if out is None:
            out = torch.empty((rows, Cn), dtype=out_dtype, device=x.device)
        s = torch.empty((rows, Cn), dtype=torch.float32, device=x.device)
        L.check(self.lib.exg_add_layernorm_fwd(_p(x), _dt(x.dtype), _p(resid), _p(s), _p(gamma), _p(beta), _p(out),
                                               _dt(out.dtype), _rm(out_map, rows), rows, Cn, float(eps), _stream()),
                "exg_add_layernorm_fwd")
        return out, s

    def layernorm_bwd(self, dy1, x, gamma, *, map1=None, dy2=None, map2=None, dres=None, dxT_dtype=None,
                      want_dx=True, dg_out=None, dxsum_out=None):
        """Backward of y = LN(x)*gamma+beta with dy = dy1[map1] (+ dy2[map2]).
        Returns (dx fp32 [rows,C] = dres + LN'(dy), dxT copy or None, dgamma fp32 [C], dbeta fp32 [C]).
        dxsum_out (fp32 [C], optional): the column sums of dx are ACCUMULATED into it (bias-gradient sink)."""
        _cuda(dy1, x, gamma, dy2, dres, dxsum_out)
        rows, Cn = x.shape
        dx = torch.empty((rows, Cn), dtype=torch.float32, device=x.device) if want_dx else None
        dxT = torch.empty((rows, Cn), dtype=dxT_dtype, device=x.device) if dxT_dtype is not None else None
        if dg_out is not None:      # (dgamma, dbeta) gradient sinks: the kernel accumulates into them
            dg = dg_out
        else:
            dg = torch.zeros(2, Cn, dtype=torch.float32, device=x.device)
        L.check(self.lib.exg_layernorm_bwd(_p(dy1), _dt(dy1.dtype), _rm(map1, rows), _p(dy2),
                                           _dt(dy2.dtype) if dy2 is not None else 0, _rm(map2, rows), _p(x),
                                           _dt(x.dtype), _p(gamma), None, None, _p(dres), _p(dx), _p(dxT),
                                           _dt(dxT.dtype) if dxT is not None else 0, _p(dg[0]), _p(dg[1]),
                                           _p(dxsum_out), rows, Cn, _stream()), "exg_layernorm_bwd")
        return dx, dxT, dg[0], dg[1]

    def add_rows(self, x, rows, *, map_x=None, x2=None, map_x2=None, add=None, add_period=0, out=None,
                 out_dtype=torch.float32, map_out=None, out2_dtype=None):
        """out[map_out(r)] = x[map_x(r)] (+ x2[map_x2(r)]) (+ add[r % add_period]) for r < rows; x, x2, add fp32."""
        _cuda(x, x2, add, out)
        Cn = x.shape[-1]
        if out is None:
            out = torch.empty((rows, Cn), dtype=out_dtype, device=x.device)
        out2 = torch.empty((rows, Cn), dtype=out2_dtype, device=x.device) if out2_dtype is not None else None
        L.check(self.lib.exg_add_rows(_p(x), _rm(map_x, rows), _p(x2), _rm(map_x2, rows), _p(add),
                                      int(add_period), _p(out), _dt(out.dtype), _rm(map_out, rows), _p(out2),
                                      _dt(out2.dtype) if out2 is not None else 0, rows, Cn, _stream()),
                "exg_add_rows")
        return out, out2

    def l2norm_fwd(self, x, rows, *, map_x=None, out_dtype=None):
        """y[r] = x[map_x(r)] / ||.||  ->  (y [rows,C], inv_norm fp32 [rows])."""
        _cuda(x)
        Cn = x.shape[-1]
        y = torch.empty((rows, Cn), dtype=out_dtype or x.dtype, device=x.device)
        inv = torch.empty(rows, dtype=torch.float32, device=x.device)
        L.check(self.lib.exg_l2norm_fwd(_p(x), _dt(x.dtype), _rm(map_x, rows), _p(y), _dt(y.dtype), _p(inv), rows,
                                        Cn, _stream()), "exg_l2norm_fwd")
        return y, inv

    def l2norm_bwd(self, dy, y, inv, dx, *, map_dx=None, accumulate=False):
        """dx[map_dx(r)] (+)= (dy[r] - y[r] <dy[r],y[r]>) * inv[r]; writes into the given dx buffer."""
        _cuda(dy, y, inv, dx)
        rows, Cn = y.shape
        L.check(self.lib.exg_l2norm_bwd(_p(dy), _dt(dy.dtype), _p(y), _dt(y.dtype), _p(inv), _p(dx), _dt(dx.dtype),
                                        _rm(map_dx, rows), int(accumulate), rows, Cn, _stream()), "exg_l2norm_bwd")
        return dx

    def reserve_sms(self, n):
        """Persistent kernels launched from now on leave n SMs to other streams (NCCL); returns the previous value."""
        return int(self.lib.exg_reserve_sms(int(n)))

    def colsum(self, x, out=None):
        """fp32 [N] = sum over rows of x [M,N]; with `out` the sums are ACCUMULATED into it (gradient sinks)."""
        _cuda(x, out)
        M, N = x.shape
        acc = out is not None
        if out is None:
            out = torch.empty(N, dtype=torch.float32, device=x.device)
        L.check(self.lib.exg_colsum(_p(x), _dt(x.dtype), M, N, _rows2d(x), _p(out), int(acc), _stream()), "exg_colsum")
        return out

    def wcolsum(self, x, row_w, out=None):
        """fp32 [N] = sum_m row_w[m] * x[m, :]; with `out` the sums are ACCUMULATED into it."""
        _cuda(x, row_w, out)
        M, N = x.shape
        acc = out is not None
        if out is None:
            out = torch.empty(N, dtype=torch.float32, device=x.device)
        L.check(self.lib.exg_wcolsum(_p(x), _dt(x.dtype), _p(row_w), M, N, _rows2d(x), _p(out), int(acc), _stream()),
                "exg_wcolsum")
        return out

    def rowdot(self, x, w, bias=None):
        """fp32 [rows] = x [rows, C] @ w [C] (+ bias[0]); w, bias fp32."""
        _cuda(x, w, bias)
        rows, Cn = x.shape
        y = torch.empty(rows, dtype=torch.float32, device=x.device)
        L.check(self.lib.exg_rowdot(_p(x), _dt(x.dtype), _p(w), _p(bias), _p(y), rows, Cn, _stream()), "exg_rowdot")
        return y

    def outer_rows(self, g, w, out_dtype):
        """[rows, C] (out_dtype) = g [rows] (fp32) outer w [C] (fp32)."""
        _cuda(g, w)
        out = torch.empty((g.numel(), w.numel()), dtype=out_dtype, device=g.device)
        L.check(self.lib.exg_outer_rows(_p(g), _p(w), _p(out), _dt(out_dtype), g.numel(), w.numel(), _stream()),
                "exg_outer_rows")
        return out

    def cast(self, src, dtype, out=None):
        _cuda(src)
        if out is None:
            out = torch.empty(src.shape, dtype=dtype, device=src.device)
        L.check(self.lib.exg_cast(_p(src), _dt(src.dtype), _p(out), _dt(out.dtype), src.numel(), _stream()), "exg_cast")
        return out

    def stage_mean(self, x, *, outT_dtype=None):
        """x [S, ...] contiguous -> (fp32 mean over S, optional T copy)."""
        _cuda(x)
        S, n = x.shape[0], x[0].numel()
        out = torch.empty(x.shape[1:], dtype=torch.float32, device=x.device)
        outT = torch.empty(x.shape[1:], dtype=outT_dtype, device=x.device) if outT_dtype is not None else None
        L.check(self.lib.exg_stage_mean(_p(x), _dt(x.dtype), x.stride(0), S, _p(out), _p(outT),
                                        _dt(outT.dtype) if outT is not None else 0, n, _stream()), "exg_stage_mean")
        return out, outT

    def bcast_rows(self, g, S, alpha, out_dtype):
        """out[s] = alpha * g for s < S (backward of stage_mean); g fp32 [...]. Returns [S, ...] of out_dtype."""
        _cuda(g)
        out = torch.empty((S,) + tuple(g.shape), dtype=out_dtype, device=g.device)
        L.check(self.lib.exg_bcast_rows(_p(g), float(alpha), _p(out), _dt(out.dtype), g.numel(), S, g.numel(),
                                        _stream()), "exg_bcast_rows")
        return out

    # ------------------------------------------------------------------ ExoGround losses
    def xent_diag_fwd(self, logits, period):
        """logits [rows, cols]; target column of row r is r % period.  Returns fp32 (loss_rows, lse_rows)."""
        _cuda(logits)
        rows, cols = logits.shape
        out = torch.empty((2, rows), dtype=torch.float32, device=logits.device)
        L.check(self.lib.exg_xent_diag_fwd(_p(logits), _dt(logits.dtype), _p(out[0]), _p(out[1]), rows, cols,
                                           int(period), _stream()), "exg_xent_diag_fwd")
        return out[0], out[1]

    def xent_diag_bwd(self, logits, lse, scale, period, g_dtype=None):
        _cuda(logits, lse)
        rows, cols = logits.shape
        G = torch.empty(logits.shape, dtype=g_dtype or logits.dtype, device=logits.device)
        L.check(self.lib.exg_xent_diag_bwd(_p(logits), _dt(logits.dtype), _p(lse), float(scale), _p(G), _dt(G.dtype),
                                           rows, cols, int(period), _stream()), "exg_xent_diag_bwd")
        return G

    def xview_nce_fwd(self, out_feat, ego, pos_idx, neg_idx, same_feat, same_idx, only_same, inv_temp):
        """out_feat [B,T,D] (T dtype), ego fp32 [B,V,T,D], idx int64 [B,T]; same_* optional.
        Returns (loss, l1, pos_cos, neg_cos) fp32 [B,T] and the stats workspace [B,T,8]."""
        _cuda(out_feat, ego, pos_idx, neg_idx, same_feat, same_idx)
        B, T, D = out_feat.shape
        V = ego.shape[1]
        res = torch.empty((4, B, T), dtype=torch.float32, device=out_feat.device)
        stats = torch.empty((B, T, 8), dtype=torch.float32, device=out_feat.device)
        L.check(self.lib.exg_xview_nce_fwd(_p(out_feat), _dt(out_feat.dtype), _p(ego), _p(pos_idx), _p(neg_idx),
                                           _p(same_feat), _p(same_idx), int(bool(only_same)), B, V, T, D,
                                           float(inv_temp), _p(res[0]), _p(res[1]), _p(res[2]), _p(res[3]),
                                           _p(stats), _stream()), "exg_xview_nce_fwd")
        return res[0], res[1], res[2], res[3], stats

    def xview_nce_bwd(self, out_feat, ego, pos_idx, neg_idx, same_feat, same_idx, only_same, inv_temp, stats, gl):
        _cuda(out_feat, ego, pos_idx, neg_idx, same_feat, same_idx, stats, gl)
        B, T, D = out_feat.shape
        V = ego.shape[1]
        go = torch.empty_like(out_feat)
        L.check(self.lib.exg_xview_nce_bwd(_p(out_feat), _dt(out_feat.dtype), _p(ego), _p(pos_idx), _p(neg_idx),
                                           _p(same_feat), _p(same_idx), int(bool(only_same)), B, V, T, D,
                                           float(inv_temp), _p(stats), _p(gl), _p(go), _dt(go.dtype), _stream()),
                "exg_xview_nce_bwd")
        return go

    def ground_loss(self, preds, gt_a, gt_b, text_pad, use_cd, eps, thresholds):
        """preds fp32 [M,2] contiguous; gt fp32 [M]; text_pad uint8 [M].  Returns (stats [4+n], iou [M],
        dpreds [M,2]); stats = (mean|d0|, mean|d1|, mean IoU, n_valid, counts(IoU > th)...)."""
        _cuda(preds, gt_a, gt_b, text_pad)
        M = preds.shape[0]
        n = len(thresholds)
        th = (C.c_float * max(n, 1))(*[float(t) for t in thresholds])
        stats = torch.empty(4 + n, dtype=torch.float32, device=preds.device)
        iou = torch.empty(M, dtype=torch.float32, device=preds.device)
        dp = torch.empty((M, 2), dtype=torch.float32, device=preds.device)
        L.check(self.lib.exg_ground_loss(_p(preds), _p(gt_a), _p(gt_b), _p(text_pad), M, int(bool(use_cd)),
                                         float(eps), th, n, _p(stats), _p(iou), _p(dp), _stream()), "exg_ground_loss")
        return stats, iou, dp

    # ------------------------------------------------------------------ attention
    def attn_fwd(self, q, k, v, key_pad, B, H, Sq, Sk, scale):
        """q [B*Sq, H*64] / k,v [B*Sk, H*64] row-strided views (T); key_pad uint8 [B,Sk] or None.
        Returns (o [B*Sq, H*64] T, lse fp32 [B,H,Sq])."""
        _cuda(q, k, v, key_pad)
        d = 64
        o = torch.empty((B * Sq, H * d), dtype=q.dtype, device=q.device)
        lse = torch.empty((B, H, Sq), dtype=torch.float32, device=q.device)
        L.check(self.lib.exg_attn_fwd(_p(q), _p(k), _p(v), _dt(q.dtype), _rows2d(q), _rows2d(k), _rows2d(v),
                                      _p(key_pad), _p(o), o.stride(0), _p(lse), B, H, Sq, Sk, d, float(scale),
                                      _stream()), "exg_attn_fwd")
        return o, lse

    def attn_bwd(self, q, k, v, o, do, lse, key_pad, dq, dk, dv, B, H, Sq, Sk, scale):
        """Writes dq/dk/dv (views with the same row strides as q/k/v)."""
        _cuda(q, k, v, o, do, lse, key_pad, dq, dk, dv)
        if (_rows2d(dq), _rows2d(dk), _rows2d(dv)) != (_rows2d(q), _rows2d(k), _rows2d(v)) or _rows2d(do) != _rows2d(o):
            raise ValueError("attn_bwd: gradient views must use the strides of their primals")
        delta = torch.empty((self.lib.exg_attn_bwd_workspace(B, H, Sq, Sk, 64),), dtype=torch.float32, device=q.device)
        L.check(self.lib.exg_attn_bwd(_p(q), _p(k), _p(v), _p(o), _p(do), _dt(q.dtype), _rows2d(q), _rows2d(k),
                                      _rows2d(v), _rows2d(o), _p(key_pad), _p(lse), _p(delta), _p(dq), _p(dk),
                                      _p(dv), B, H, Sq, Sk, 64, float(scale), _stream()), "exg_attn_bwd")

    # ------------------------------------------------------------------ MIL-NCE
    def nce_pack_targets(self, tgt):
        """tgt uint8 [B,N,T] -> int32 [B*T, ceil(N/32)] bit rows (bit n of row (b,t) = tgt[b,n,t])."""
        _cuda(tgt)
        B, N, T = tgt.shape
        bits = torch.empty((B * T, (N + 31) // 32), dtype=torch.int32, device=tgt.device)
        L.check(self.lib.exg_nce_pack_targets(_p(tgt), _p(bits), B, T, N, _stream()), "exg_nce_pack_targets")
        return bits

    def nce_fwd(self, logits, tgt, col_valid, B, T, N, inv_temp, tgt_bits=None, own_row_mask=None, Bc=None,
                clip0=0):
        """logits [S, B*T, Bc*N] contiguous (T); tgt uint8 [B,N,T]; col_valid uint8 [Bc*N]; tgt_bits: optional
        nce_pack_targets(tgt) (packed here when omitted).  Bc / clip0: column clips and the column clip owned by
        row clip 0 (cross-GPU negatives: Bc = world * B, clip0 = rank * B; default Bc = B, clip0 = 0).
        Returns fp32 (row_all [S,R], row_pos [S,R], col_all [S,Cc], col_pos [S,Cc]) exp-sums."""
        _cuda(logits, tgt, col_valid, tgt_bits, own_row_mask)
        if tgt_bits is None:
            tgt_bits = self.nce_pack_targets(tgt)
        S = logits.shape[0]
        Bc = B if Bc is None else Bc
        assert logits.shape[1] == B * T and logits.shape[2] == Bc * N and col_valid.numel() == Bc * N
        rows = torch.empty((2, S, B * T), dtype=torch.float32, device=logits.device)
        cols = torch.zeros((2, S, Bc * N), dtype=torch.float32, device=logits.device)
        L.check(self.lib.exg_nce_fwd(_p(logits), _dt(logits.dtype), _p(tgt), _p(tgt_bits), _p(col_valid), _p(own_row_mask),
                                     _p(rows[0]), _p(rows[1]),
                                     _p(cols[0]), _p(cols[1]), S, B, T, N, Bc, clip0, float(inv_temp), _stream()),
                "exg_nce_fwd")
        return rows[0], rows[1], cols[0], cols[1]

    def nce_bwd(self, logits, tgt, col_valid, sums, wr, wc, B, T, N, inv_temp, g_dtype=None, tgt_bits=None,
                own_row_mask=None, Bc=None, clip0=0):
        """G [S,R,Cc] = d(sum_r wr*(log row_all-log row_pos) + sum_c wc*(log col_all-log col_pos)) / d logits."""
        _cuda(logits, tgt, col_valid, wr, wc, tgt_bits, own_row_mask)
        if tgt_bits is None:
            tgt_bits = self.nce_pack_targets(tgt)
        S = logits.shape[0]
        G = torch.empty(logits.shape, dtype=g_dtype or logits.dtype, device=logits.device)
        ra, rp, ca, cp = sums
        L.check(self.lib.exg_nce_bwd(_p(logits), _dt(logits.dtype), _p(tgt), _p(tgt_bits), _p(col_valid), _p(own_row_mask),
                                     _p(ra), _p(rp), _p(ca),
                                     _p(cp), _p(wr), _p(wc), _p(G), _dt(G.dtype), S, B, T, N, B if Bc is None else Bc,
                                     clip0, float(inv_temp),
                                     _stream()), "exg_nce_bwd")
        return G

    def tan_loss_tail(self, tgt, valid, sums_d, sums_j, own_max_d, own_max_j, la, abs_c, B, T, N, inv_temp,
                      loss_threshold, use_head, nce_weight, tgt_bits=None):
        """Scalar tail of the TAN loss (train/loss.py:251-360) in two launches (csrc/loss_tail.cu).
        tgt uint8 [B,N,T] (AND-ed with the valid sentences), valid uint8 [B*N]; sums_x = the four exp-sum views
        returned by nce_fwd (rows [2,S,R] / cols [2,S,Cc] storage); own_max_x fp32 [B*N]; la fp32 [B*N] or None;
        abs_c fp32 [B*N] or None.  Returns (out fp32 [16], w fp32 [5, L]) with L = padded max(R, Cc): rows of w =
        wr_dual, wc_dual, wr_joint, wc_joint, d loss / d la."""
        _cuda(tgt, valid, own_max_d, own_max_j, la, abs_c, tgt_bits, *sums_d, *sums_j)
        if tgt_bits is None:
            tgt_bits = self.nce_pack_targets(tgt)
        E, D = sums_d[0].shape[0], sums_j[0].shape[0]
        R, Cc = B * T, B * N
        dev = tgt.device
        for sm, S in ((sums_d, E), (sums_j, D)):
            ra, rp, ca, cp = sm
            assert rp.data_ptr() == ra.data_ptr() + 4 * S * R and cp.data_ptr() == ca.data_ptr() + 4 * S * Cc, \
                "tan_loss_tail expects the [2,S,R] / [2,S,Cc] buffers of nce_fwd"
        Lp = (max(R, Cc) + 63) // 64 * 64
        v = torch.empty((4, Lp), dtype=torch.float32, device=dev)
        w = torch.zeros((5, Lp), dtype=torch.float32, device=dev)
        out = torch.zeros(16, dtype=torch.float32, device=dev)
        L.check(self.lib.exg_nce_logdiff(_p(sums_d[0]), _p(sums_d[2]), E, _p(sums_j[0]), _p(sums_j[2]), D, R, Cc,
                                         _p(v[0]), _p(v[1]), _p(v[2]), _p(v[3]), _stream()), "exg_nce_logdiff")
        nws = int(self.lib.exg_tan_loss_tail_workspace(B, T, N))
        ws = torch.empty(nws, dtype=torch.uint8, device=dev)
        L.check(self.lib.exg_tan_loss_tail(_p(tgt_bits), _p(valid), _p(v[0]), _p(v[1]), _p(v[2]), _p(v[3]),
                                           _p(own_max_d), _p(own_max_j), _p(la), _p(abs_c), B, T, N, E, D,
                                           float(inv_temp), float(loss_threshold), int(bool(use_head)),
                                           float(nce_weight), _p(w[0]), _p(w[1]), _p(w[2]), _p(w[3]), _p(w[4]),
                                           _p(out), _p(ws), nws, _stream()), "exg_tan_loss_tail")
        return out, w

    def own_max(self, logits, stage, B, T, N, Bc=None, clip0=0):
        _cuda(logits)
        out = torch.empty((B, N), dtype=torch.float32, device=logits.device)
        L.check(self.lib.exg_own_max(_p(logits), _dt(logits.dtype), _p(out), int(stage), B, T, N,
                                     B if Bc is None else Bc, clip0, _stream()),
                "exg_own_max")
        return out

    # ------------------------------------------------------------------ alignment evaluation
    def align_finish(self, sj, sd, win, K, vlen):
        """sj / sd fp32 [W, T, N] window similarities (already / 0.07); win int32 [W, 4] = (first second, seconds,
        first sentence, last sentence).  Returns (sim [K, vlen], prob [K, vlen], pred int64 [K], row_max [K])."""
        _cuda(sj, sd, win)
        W, T, N = sj.shape
        dev = sj.device
        sim = torch.empty((K, vlen), dtype=torch.float32, device=dev)
        prob = torch.empty((K, vlen), dtype=torch.float32, device=dev)
        pred = torch.empty(K, dtype=torch.int64, device=dev)
        mx = torch.empty(K, dtype=torch.float32, device=dev)
        L.check(self.lib.exg_align_finish(_p(sj), _p(sd), _p(win), W, T, N, K, vlen, _p(sim), _p(prob), _p(pred),
                                          _p(mx), _stream()), "exg_align_finish")
        return sim, prob, pred, mx

    # ------------------------------------------------------------------ optimizer
    def adamw(self, param, grad, m, v, shadow, lr, beta1, beta2, eps, wd, step, grad_scale=1.0):
        _cuda(param, grad, m, v, shadow)
        L.check(self.lib.exg_adamw(_p(param), _p(grad), _p(m), _p(v), _p(shadow), param.numel(), lr, beta1, beta2,
                                   eps, wd, int(step), float(grad_scale), _stream()), "exg_adamw")

    def adamw_dev(self, param, grad, m, v, shadow, hyper, beta1, beta2, eps, wd, grad_scale=1.0):
        """adamw with hyper = device fp32 [3] {lr, 1 - beta1^step, sqrt(1 - beta2^step)} (CUDA-graph capturable)."""
        _cuda(param, grad, m, v, shadow, hyper)
        L.check(self.lib.exg_adamw_dev(_p(param), _p(grad), _p(m), _p(v), _p(shadow), param.numel(), _p(hyper),
                                       beta1, beta2, eps, wd, float(grad_scale), _stream()), "exg_adamw_dev")

    def grad_sumsq(self, grad, blk2seg, n_seg, out=None):
        """fp32 [n_seg]: squared gradient norm of every parameter of the flat buffer `grad` (fp32, numel % 64 == 0);
        blk2seg int32 [numel / 64] = parameter index of each 64-element block."""
        _cuda(grad, blk2seg, out)
        if out is None:
            out = torch.empty(n_seg, dtype=torch.float32, device=grad.device)
        L.check(self.lib.exg_grad_sumsq(_p(grad), _p(blk2seg), grad.numel(), _p(out), int(n_seg), _stream()),
                "exg_grad_sumsq")
        return out

    def grad_clip_scale(self, grad, blk2seg, sumsq, clip, grad_scale=1.0):
        """In place: every parameter's gradient *= clip / (||g|| * grad_scale + 1e-6) where that is < 1."""
        _cuda(grad, blk2seg, sumsq)
        L.check(self.lib.exg_grad_clip_scale(_p(grad), _p(blk2seg), grad.numel(), _p(sumsq), float(clip),
                                             float(grad_scale), _stream()), "exg_grad_clip_scale")

    def adamw_clip(self, param, grad, m, v, shadow, lr, beta1, beta2, eps, wd, step, blk2seg, sumsq, clip,
                   grad_scale=1.0, hyper=None):
        """adamw with the per-parameter clip factor folded into the gradient read (blk2seg indexed from this
        segment's first 64-element block)."""
        _cuda(param, grad, m, v, shadow, blk2seg, sumsq, hyper)
        L.check(self.lib.exg_adamw_clip(_p(param), _p(grad), _p(m), _p(v), _p(shadow), param.numel(), _p(hyper),
                                        float(lr), beta1, beta2, eps, wd, int(step), float(grad_scale), _p(blk2seg),
                                        _p(sumsq), float(clip), _stream()), "exg_adamw_clip")

    def ema(self, target, online, m):
        _cuda(target, online)
        L.check(self.lib.exg_ema(_p(target), _p(online), target.numel(), float(m), _stream()), "exg_ema")
