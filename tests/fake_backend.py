"""Plain-torch restatement of the op specs of exoground_b200/ops/backend.py.

TEST INFRASTRUCTURE: (1) the per-kernel reference the GPU unit tests compare libexg.so against, and
(2) injected with ops.set_backend() by the CPU tests so the host logic (buffer plumbing, manual
backward chaining, row maps) can be validated against the oracle without a GPU.  Never shipped.
All math is done in fp32 and rounded to the requested output dtype, like the kernels do.
"""
import torch

from exoground_b200.ops import lib as L


def _map_idx(rm, rows, device):
    r = torch.arange(rows, device=device)
    if rm is None:
        return r
    g, gs, off = rm
    return (r // g) * gs + (r % g) + off


def _qgelu(v):
    return v * torch.sigmoid(1.702 * v)


def _qgelu_grad(u):
    s = torch.sigmoid(1.702 * u)
    return s + 1.702 * u * s * (1 - s)


class TorchBackend:
    name = "fake-torch"

    def gemm(self, A, B, *, transA=False, transB=False, bias=None, epilogue=L.EPI_NONE, aux=None,
             out_dtype=None, alpha=1.0, out=None, want_preact=False, impl=0):
        a = A.float().transpose(-1, -2) if transA else A.float()
        b = B.float() if transB else B.float().transpose(-1, -2)
        v = alpha * (a @ b)
        if bias is not None:
            v = v + bias.float()
        pre = v
        if epilogue == L.EPI_QGELU:
            o = _qgelu(v)
        elif epilogue == L.EPI_RESID:
            o = aux.float() + v
        elif epilogue == L.EPI_QGELU_BWD:
            o = v * _qgelu_grad(aux.float())
        elif epilogue == L.EPI_RELU:
            o = v.clamp(min=0)
        elif epilogue == L.EPI_RELU_BWD:
            o = v * (aux.float() > 0)
        else:
            o = v
        out_dtype = out_dtype or A.dtype
        if out is not None:
            out.copy_(o.to(out.dtype))
        else:
            out = o.to(out_dtype)
        if want_preact:
            return out, pre.to(out.dtype)
        return out

    def layernorm_fwd(self, x, gamma, beta, *, out_dtype, out=None, out_map=None, y2_dtype=None, add=None,
                      add_period=0, eps=1e-5):
        rows, C = x.shape
        xf = x.float()
        mu = xf.mean(-1, keepdim=True)
        var = ((xf - mu) ** 2).mean(-1, keepdim=True)
        y = (xf - mu) * torch.rsqrt(var + eps) * gamma + beta
        if add is not None:
            y = y + add[torch.arange(rows, device=x.device) % add_period]
        if out is None:
            out = torch.empty(rows, C, dtype=out_dtype, device=x.device)
        out.view(-1, C)[_map_idx(out_map, rows, x.device)] = y.to(out.dtype)
        y2 = y.to(y2_dtype) if y2_dtype is not None else None
        return out, y2

    def add_layernorm_fwd(self, x, resid, gamma, beta, *, out_dtype, out=None, out_map=None, eps=1e-5):
        s = resid.float() + x.float()
        y, _ = self.layernorm_fwd(s, gamma, beta, out_dtype=out_dtype, out=out, out_map=out_map, eps=eps)
        return y, s

    def layernorm_bwd(self, dy1, x, gamma, *, map1=None, dy2=None, map2=None, dres=None, dxT_dtype=None,
                      want_dx=True, dg_out=None, dxsum_out=None):
        rows, C = x.shape
        dy = dy1.view(-1, C)[_map_idx(map1, rows, x.device)].float()
        if dy2 is not None:
            dy = dy + dy2.view(-1, C)[_map_idx(map2, rows, x.device)].float()
        xf = x.float()
        mu = xf.mean(-1, keepdim=True)
        var = ((xf - mu) ** 2).mean(-1, keepdim=True)
        rstd = torch.rsqrt(var + 1e-5)
        xhat = (xf - mu) * rstd
        dxh = dy * gamma
        dx = (dxh - dxh.mean(-1, keepdim=True) - xhat * (dxh * xhat).mean(-1, keepdim=True)) * rstd
        if dres is not None:
            dx = dx + dres
        dxT = dx.to(dxT_dtype) if dxT_dtype is not None else None
        if dxsum_out is not None:
            dxsum_out.add_(dx.sum(0))
        dgam, dbet = (dy * xhat).sum(0), dy.sum(0)
        if dg_out is not None:
            dg_out[0].add_(dgam); dg_out[1].add_(dbet)
            dgam, dbet = dg_out[0], dg_out[1]
        return (dx if want_dx else None), dxT, dgam, dbet

    def add_rows(self, x, rows, *, map_x=None, x2=None, map_x2=None, add=None, add_period=0, out=None,
                 out_dtype=torch.float32, map_out=None, out2_dtype=None):
        C = x.shape[-1]
        v = x.view(-1, C)[_map_idx(map_x, rows, x.device)].float()
        if x2 is not None:
            v = v + x2.view(-1, C)[_map_idx(map_x2, rows, x.device)].float()
        if add is not None:
            v = v + add[torch.arange(rows, device=x.device) % add_period]
        if out is None:
            out = torch.empty(rows, C, dtype=out_dtype, device=x.device)
        out.view(-1, C)[_map_idx(map_out, rows, x.device)] = v.to(out.dtype)
        out2 = v.to(out2_dtype) if out2_dtype is not None else None
        return out, out2

    def l2norm_fwd(self, x, rows, *, map_x=None, out_dtype=None):
        C = x.shape[-1]
        v = x.view(-1, C)[_map_idx(map_x, rows, x.device)].float()
        inv = 1.0 / v.norm(dim=-1)
        return (v * inv[:, None]).to(out_dtype or x.dtype), inv

    def l2norm_bwd(self, dy, y, inv, dx, *, map_dx=None, accumulate=False):
        rows, C = y.shape
        g, yy = dy.float(), y.float()
        d = (g - yy * (g * yy).sum(-1, keepdim=True)) * inv[:, None]
        idx = _map_idx(map_dx, rows, y.device)
        dv = dx.view(-1, C)
        if accumulate:
            d = d + dv[idx].float()
        dv[idx] = d.to(dx.dtype)
        return dx

    def reserve_sms(self, n):
        return 0

    def colsum(self, x, out=None):
        if out is not None:
            out += x.float().sum(0)
            return out
        return x.float().sum(0)

    def wcolsum(self, x, row_w, out=None):
        r = (x.float() * row_w[:, None]).sum(0)
        if out is not None:
            out += r
            return out
        return r

    def rowdot(self, x, w, bias=None):
        y = x.float() @ w.float()
        return y + bias[0] if bias is not None else y

    def outer_rows(self, g, w, out_dtype):
        return (g.float()[:, None] * w.float()[None, :]).to(out_dtype)

    def cast(self, src, dtype, out=None):
        if out is None:
            return src.to(dtype)
        out.copy_(src.to(out.dtype))
        return out

    def stage_mean(self, x, *, outT_dtype=None):
        m = x.float().mean(0)
        return m, (m.to(outT_dtype) if outT_dtype is not None else None)

    def bcast_rows(self, g, S, alpha, out_dtype):
        return (g.float() * alpha).to(out_dtype).unsqueeze(0).repeat((S,) + (1,) * g.dim())

    def xent_diag_fwd(self, logits, period):
        x = logits.float()
        rows = x.shape[0]
        lse = torch.logsumexp(x, dim=1)
        tgt = torch.arange(rows, device=x.device) % period
        return lse - x[torch.arange(rows, device=x.device), tgt], lse

    def xent_diag_bwd(self, logits, lse, scale, period, g_dtype=None):
        x = logits.float()
        rows = x.shape[0]
        g = torch.exp(x - lse[:, None])
        tgt = torch.arange(rows, device=x.device) % period
        g[torch.arange(rows, device=x.device), tgt] -= 1.0
        return (g * scale).to(g_dtype or logits.dtype)

    @staticmethod
    def _xview_parts(out_feat, ego, pos_idx, neg_idx, same_feat, same_idx):
        B, T, D = out_feat.shape
        bi = torch.arange(B, device=ego.device)[:, None]
        ti = torch.arange(T, device=ego.device)[None, :]
        o = out_feat.float()
        ks = [ego[bi, pos_idx, ti], ego[bi, neg_idx, ti]]
        if same_feat is not None:
            ks.append(same_feat[bi, same_idx])
        nrm = lambda x: x / x.norm(dim=2, keepdim=True).clamp(min=1e-12)
        return o, nrm(o), [nrm(k.float()) for k in ks]

    def xview_nce_fwd(self, out_feat, ego, pos_idx, neg_idx, same_feat, same_idx, only_same, inv_temp):
        o, oh, kh = self._xview_parts(out_feat, ego, pos_idx, neg_idx, same_feat, same_idx)
        c = [(oh * k).sum(2) for k in kh]
        use = [0, 2] if (same_feat is not None and only_same) else list(range(len(kh)))
        sims = torch.stack([c[i] * inv_temp for i in use], 2)
        loss = -torch.log_softmax(sims, dim=2)[:, :, 0]
        l1 = (oh - kh[0]).abs().mean(2)
        return loss, l1, c[0], c[1], None

    def xview_nce_bwd(self, out_feat, ego, pos_idx, neg_idx, same_feat, same_idx, only_same, inv_temp, stats, gl):
        with torch.enable_grad():
            o = out_feat.detach().float().requires_grad_(True)
            loss = self.xview_nce_fwd(o, ego, pos_idx, neg_idx, same_feat, same_idx, only_same, inv_temp)[0]
            (g,) = torch.autograd.grad((loss * gl).sum(), o)
        return g.to(out_feat.dtype)

    def ground_loss(self, preds, gt_a, gt_b, text_pad, use_cd, eps, thresholds):
        with torch.enable_grad():
            p = preds.detach().float().requires_grad_(True)
            v = ~text_pad.bool()
            a, b = p[:, 0], p[:, 1]
            if use_cd:
                ps, pe, gs, ge = a - b / 2, a + b / 2, gt_a - gt_b / 2, gt_a + gt_b / 2
            else:
                ps, pe, gs, ge = a, b, gt_a, gt_b
            inter = (torch.minimum(pe, ge) - torch.maximum(ps, gs)).clamp(min=0)
            union = torch.maximum(pe, ge) - torch.minimum(ps, gs)
            iou = torch.where(v, inter / (union + eps), torch.zeros_like(inter))
            n = v.float().sum()
            l1a = ((a - gt_a).abs() * v).sum() / n
            l1b = ((b - gt_b).abs() * v).sum() / n
            miou = iou.sum() / n
            loss = (1 - miou) + (l1a + l1b if use_cd else (l1a + l1b) / 2)
            (dp,) = torch.autograd.grad(loss, p)
        cnt = [((iou > t) & v).float().sum() for t in thresholds]
        stats = torch.stack([l1a.detach(), l1b.detach(), miou.detach(), n] + cnt)
        return stats, iou.detach(), dp

    def attn_fwd(self, q, k, v, key_pad, B, H, Sq, Sk, scale):
        d = 64
        qf = q.float().reshape(B, Sq, H, d).transpose(1, 2)
        kf = k.float().reshape(B, Sk, H, d).transpose(1, 2)
        vf = v.float().reshape(B, Sk, H, d).transpose(1, 2)
        s = qf @ kf.transpose(-1, -2) * scale
        if key_pad is not None:
            s = s.masked_fill(key_pad.bool()[:, None, None, :], float("-inf"))
        lse = torch.logsumexp(s, dim=-1)
        p = torch.exp(s - lse[..., None])
        o = (p @ vf).transpose(1, 2).reshape(B * Sq, H * d)
        return o.to(q.dtype), lse

    def attn_bwd(self, q, k, v, o, do, lse, key_pad, dq, dk, dv, B, H, Sq, Sk, scale):
        d = 64
        qf = q.float().reshape(B, Sq, H, d).transpose(1, 2)
        kf = k.float().reshape(B, Sk, H, d).transpose(1, 2)
        vf = v.float().reshape(B, Sk, H, d).transpose(1, 2)
        dof = do.float().reshape(B, Sq, H, d).transpose(1, 2)
        of = o.float().reshape(B, Sq, H, d).transpose(1, 2)
        s = qf @ kf.transpose(-1, -2) * scale
        if key_pad is not None:
            s = s.masked_fill(key_pad.bool()[:, None, None, :], float("-inf"))
        p = torch.exp(s - lse[..., None])
        delta = (dof * of).sum(-1, keepdim=True)
        dp = dof @ vf.transpose(-1, -2)
        ds = p * (dp - delta) * scale
        dq.copy_((ds @ kf).transpose(1, 2).reshape(B * Sq, H * d).to(dq.dtype))
        dk.copy_((ds.transpose(-1, -2) @ qf).transpose(1, 2).reshape(B * Sk, H * d).to(dk.dtype))
        dv.copy_((p.transpose(-1, -2) @ dof).transpose(1, 2).reshape(B * Sk, H * d).to(dv.dtype))

    @staticmethod
    def _own(B, T, N, Bc, clip0, device):
        """bool [B*T, Bc*N]: column (b', n) belongs to the clip of row (a, t), i.e. b' == a + clip0."""
        rc = torch.arange(B, device=device).repeat_interleave(T) + clip0
        cc = torch.arange(Bc, device=device).repeat_interleave(N)
        return rc[:, None] == cc[None, :]

    def _nce_common(self, logits, tgt, col_valid, B, T, N, inv_temp, own_row_mask=None, Bc=None, clip0=0):
        Bc = B if Bc is None else Bc
        e = torch.exp(logits.float() * inv_temp - inv_temp)                     # [S,R,Cc]
        valid = col_valid.bool().view(1, 1, Bc * N)
        e = torch.where(valid, e, torch.zeros_like(e))
        own = self._own(B, T, N, Bc, clip0, logits.device)
        if own_row_mask is not None:       # own-clip block of the masked rows counts as -inf
            drop = own & own_row_mask.bool().view(B * T, 1)
            e = torch.where(drop[None], torch.zeros_like(e), e)
        t_bt_n = tgt.bool().permute(0, 2, 1).reshape(B * T, N)                  # row (a,t) x sentence n
        pos = own & t_bt_n.repeat(1, Bc) & col_valid.bool().view(1, Bc * N)
        return e, pos

    def nce_pack_targets(self, tgt):
        B, N, T = tgt.shape
        W = (N + 31) // 32
        t = torch.zeros((B, W * 32, T), dtype=torch.int64, device=tgt.device)
        t[:, :N] = (tgt != 0).long()
        sh = (torch.arange(W * 32, device=tgt.device) % 32).view(1, -1, 1)
        words = (t << sh).view(B, W, 32, T).sum(2)                       # [B, W, T]
        words = torch.where(words >= 2 ** 31, words - 2 ** 32, words)     # two's complement int32
        return words.permute(0, 2, 1).reshape(B * T, W).to(torch.int32)

    def nce_fwd(self, logits, tgt, col_valid, B, T, N, inv_temp, tgt_bits=None, own_row_mask=None, Bc=None, clip0=0):
        e, pos = self._nce_common(logits, tgt, col_valid, B, T, N, inv_temp, own_row_mask, Bc, clip0)
        ep = e * pos[None]
        return e.sum(-1), ep.sum(-1), e.sum(-2), ep.sum(-2)

    def nce_bwd(self, logits, tgt, col_valid, sums, wr, wc, B, T, N, inv_temp, g_dtype=None, tgt_bits=None,
                own_row_mask=None, Bc=None, clip0=0):
        Bc = B if Bc is None else Bc
        e, pos = self._nce_common(logits, tgt, col_valid, B, T, N, inv_temp, own_row_mask, Bc, clip0)
        ra, rp, ca, cp = sums
        safe = lambda w, s: torch.where((w != 0) & (s > 0), w / s.clamp(min=1e-38), torch.zeros_like(s))
        a_r = safe(wr[None, :].expand_as(ra), ra)[..., None]
        b_r = safe(wr[None, :].expand_as(rp), rp)[..., None]
        a_c = safe(wc[None, :].expand_as(ca), ca)[:, None, :]
        b_c = safe(wc[None, :].expand_as(cp), cp)[:, None, :]
        coef = a_r + a_c - pos[None].float() * (b_r + b_c)
        G = inv_temp * e * coef
        if own_row_mask is not None:
            # positives removed by the own-row mask: where ALL positives of a weighted row (column) are gone, the
            # reference's positive LSE runs over -6e4 fills only and leaks -1/#entries through each of them
            mpos = (pos & own_row_mask.bool().view(B * T, 1))[None].float()
            n_cols = col_valid.float().sum()
            lr_ = torch.where((rp <= 0) & (wr[None, :] != 0), wr[None, :].expand_as(rp) / n_cols, torch.zeros_like(rp))
            lc_ = torch.where((cp <= 0) & (wc[None, :] != 0), wc[None, :].expand_as(cp) / float(Bc * T),
                              torch.zeros_like(cp))
            G = G - inv_temp * mpos * (lr_[..., None] + lc_[:, None, :])
        return G.to(g_dtype or logits.dtype)

    def tan_loss_tail(self, tgt, valid, sums_d, sums_j, own_max_d, own_max_j, la, abs_c, B, T, N, inv_temp,
                      loss_threshold, use_head, nce_weight, tgt_bits=None):
        """Spec of exg_nce_logdiff + exg_tan_loss_tail, written the way the reference computes it
        (train/loss.py:251-360: boolean compaction, torch.quantile, Tensor.std, F.binary_cross_entropy_with_logits)."""
        R, Cc = B * T, B * N
        tg = tgt.bool().view(B, N, T)
        vflat = valid.bool().view(-1)
        rowp, colp = tg.any(dim=1).reshape(-1), tg.any(dim=2).reshape(-1)

        def masked_mean_loss(sums, rmask, cmask):
            ra, rp, ca, cp = sums
            v = torch.log(ra) - torch.log(rp.clamp(min=1e-38))
            t = torch.log(ca.clamp(min=1e-38)) - torch.log(cp.clamp(min=1e-38))
            return (v[:, rmask].mean() + t[:, cmask].mean()) / 2

        E, D = sums_d[0].shape[0], sums_j[0].shape[0]
        ld_all, lj_all = masked_mean_loss(sums_d, rowp, colp), masked_mean_loss(sums_j, rowp, colp)
        ld, lj, rmask, cmask = ld_all, lj_all, rowp, colp
        md, mj = own_max_d.reshape(-1) * inv_temp, own_max_j.reshape(-1) * inv_temp
        if loss_threshold > 0:
            zd = (md - md[vflat].mean()) / md[vflat].std()
            zj = (mj - mj[vflat].mean()) / mj[vflat].std()
            metric = -(zd + zj)
            keep = vflat & (metric <= torch.quantile(metric[vflat], loss_threshold))
            rmask = (tg & keep.view(B, N, 1)).any(dim=1).reshape(-1)
            cmask = keep
            ld, lj = masked_mean_loss(sums_d, rmask, cmask), masked_mean_loss(sums_j, rmask, cmask)
        Lp = (max(R, Cc) + 63) // 64 * 64
        w = torch.zeros((5, Lp), dtype=torch.float32, device=tgt.device)
        out = torch.zeros(16, dtype=torch.float32, device=tgt.device)
        bce = top1 = torch.zeros((), device=tgt.device)
        if use_head:
            qd, qj = torch.quantile(md[vflat], 0.5), torch.quantile(mj[vflat], 0.5)
            lab = torch.full_like(md, 2.0)
            lab[(md > qd) & (mj > qj)] = 1.0
            lab[(md < qd) & (mj < qj)] = 0.0
            if abs_c is not None:
                lab[(abs_c < 0.2) | (abs_c > 0.8)] = 0.0
            sel = vflat & (lab != 2)
            y = lab[sel]
            x = la.detach().clone().requires_grad_(True)
            with torch.enable_grad():
                bce_t = torch.nn.functional.binary_cross_entropy_with_logits(x[sel], y, pos_weight=1 / y.mean() - 1)
                bce_t.backward()
            w[4, :Cc] = x.grad
            bce = bce_t.detach()
            top1 = ((x.detach()[sel] > 0).float() == y).float().mean()
        scale = 0.5 * (nce_weight if use_head else 1.0)
        nr, nc = rmask.float().sum(), cmask.float().sum()
        w[0, :R] = rmask.float() * scale / (2 * E * nr)
        w[1, :Cc] = cmask.float() * scale / (2 * E * nc)
        w[2, :R] = rmask.float() * scale / (2 * D * nr)
        w[3, :Cc] = cmask.float() * scale / (2 * D * nc)
        loss = (ld + lj) / 2
        if use_head:
            loss = loss * nce_weight + bce
        for i, val in enumerate((loss, ld, lj, ld_all, lj_all, (ld_all + lj_all) / 2, bce, top1, nr, nc, vflat.float().sum())):
            out[i] = val
        return out, w

    def own_max(self, logits, stage, B, T, N, Bc=None, clip0=0):
        Bc = B if Bc is None else Bc
        lg = logits[stage].float().view(B, T, Bc, N)
        return torch.stack([lg[b, :, b + clip0, :].max(0).values for b in range(B)], 0)

    def adamw(self, param, grad, m, v, shadow, lr, beta1, beta2, eps, wd, step, grad_scale=1.0):
        g = grad * grad_scale
        param.mul_(1 - lr * wd)
        m.mul_(beta1).add_(g, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
        bc1, bc2 = 1 - beta1 ** step, 1 - beta2 ** step
        param.addcdiv_(m, v.sqrt() / (bc2 ** 0.5) + eps, value=-lr / bc1)
        if shadow is not None:
            shadow.copy_(param.to(shadow.dtype))

    def adamw_dev(self, param, grad, m, v, shadow, hyper, beta1, beta2, eps, wd, grad_scale=1.0):
        lr, bc1, bc2s = (float(x) for x in hyper)
        g = grad * grad_scale
        param.mul_(1 - lr * wd)
        m.mul_(beta1).add_(g, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
        param.addcdiv_(m, v.sqrt() / bc2s + eps, value=-lr / bc1)
        if shadow is not None:
            shadow.copy_(param)

    def gemm_add_ln_ok(self, A, W):
        import os
        return (os.environ.get("EXG_GEMM_LN", "0") == "1" and A.dtype == torch.bfloat16 and W.dtype == torch.bfloat16
                and A.dim() == 2 and W.dim() == 2 and W.shape[0] == 512 and A.shape[1] % 8 == 0 and A.shape[1] >= 64)

    def gemm_add_ln(self, A, W, bias, resid, gamma, beta, *, out=None, eps=1e-5):
        s = resid + A.float() @ W.float().t() + (bias if bias is not None else 0.0)
        y = torch.nn.functional.layer_norm(s, (512,), gamma, beta, eps).to(torch.bfloat16)
        if out is not None:
            out.copy_(y)
            y = out
        return y, s

    def align_finish(self, sj, sd, win, K, vlen):
        """Spec = the reference's loop (eval_zeroshot_align.py:200-222), windows in order."""
        logits = torch.zeros(K, vlen)
        dual = torch.zeros(K, vlen)
        count = torch.zeros(K, vlen)
        for w, (s, ln, l, r) in enumerate(win.tolist()):
            n = r - l + 1
            logits[l:r + 1, s:s + ln] += sj[w, :ln, :n].t().float().cpu()
            dual[l:r + 1, s:s + ln] += sd[w, :ln, :n].t().float().cpu()
            count[l:r + 1, s:s + ln] += 1
        eps = torch.tensor(1e-5)
        sim = (logits / torch.maximum(count, eps) + dual / torch.maximum(count, eps)) / 2
        sim = sim.masked_fill(sim == 0, -6e4).to(sj.device)
        prob = sim.softmax(-1)
        return sim, prob, prob.argmax(-1), sim.max(-1).values

    def grad_sumsq(self, grad, blk2seg, n_seg, out=None):
        sq = grad.double().pow(2).view(-1, 64).sum(1)
        res = torch.zeros(n_seg, dtype=torch.float64, device=grad.device).index_add_(0, blk2seg.long(), sq).float()
        if out is not None:
            out.copy_(res)
            return out
        return res

    @staticmethod
    def _clip_coef(blk2seg, sumsq, clip, grad_scale):
        c = clip / (sumsq.sqrt() * grad_scale + 1e-6)
        return c.clamp(max=1.0)[blk2seg.long()].repeat_interleave(64)

    def grad_clip_scale(self, grad, blk2seg, sumsq, clip, grad_scale=1.0):
        grad.mul_(self._clip_coef(blk2seg, sumsq, clip, grad_scale))

    def adamw_clip(self, param, grad, m, v, shadow, lr, beta1, beta2, eps, wd, step, blk2seg, sumsq, clip,
                   grad_scale=1.0, hyper=None):
        g = grad * self._clip_coef(blk2seg, sumsq, clip, grad_scale)[:grad.numel()]
        if hyper is not None:
            self.adamw_dev(param, g, m, v, shadow, hyper, beta1, beta2, eps, wd, grad_scale)
        else:
            self.adamw(param, g, m, v, shadow, lr, beta1, beta2, eps, wd, step, grad_scale)

    def ema(self, target, online, m):
        target.mul_(m).add_(online, alpha=1 - m)
