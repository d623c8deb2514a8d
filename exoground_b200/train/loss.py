"""Drop-in replacement of the reference's train/loss.py: get_loss / get_mask_from_time / get_text_pos.

The logits-sized work (the masked bidirectional MIL-NCE of loss.py:234-278 and its gradient, the
per-text max of loss.py:283-286) runs in libexg.so kernels (exg_nce_fwd / exg_nce_bwd / exg_own_max);
the remaining arithmetic touches vectors of at most B*T or B*N elements and is written with
static-shape masked torch ops (no boolean indexing, no .item()), so the whole step stays
CUDA-graph capturable.  `logits` may come from exoground_b200's TemporalAligner (zero-copy
stage-major storage) or from the reference model (copied once into stage-major layout).
"""
import torch
import torch.nn.functional as F
from torch.nn.utils.rnn import pad_sequence

from .. import ops
from ..model.functional import NceLossFn, nce_value

INV_TEMP = 1.0 / 0.07      # train/loss.py:68-69 (logits / 0.07 for args.sim == 'cos')


def circulant(tensor, dim):
    """Circulant expansion along `dim`, new axis last: circulant([0,1,2], 0) = [[0,1,2],[2,0,1],[1,2,0]]
    (train/loss.py:12-20 / loss_egoexo4d.py:18-26, used by the self-labelling branch); out[.., i, .., j] =
    tensor[.., (j - i) mod S, ..], one index_select instead of the reference's flip / cat / unfold."""
    S = tensor.shape[dim]
    dim = dim % tensor.dim()
    ar = torch.arange(S, device=tensor.device)
    idx = (ar[None, :] - ar[:, None]) % S
    return tensor.index_select(dim, idx.reshape(-1)).unflatten(dim, (S, S)).movedim(dim + 1, -1)


def get_mask_from_time(start_list, end_list, num_timestamp, num_text, device='cuda'):
    """bool [B, num_text, T]: 1 where start <= t < end (reference train/loss.py:26-43, minus its prints).
    Built on the host from the python lists and copied once."""
    start = pad_sequence([torch.FloatTensor(i) for i in start_list], batch_first=True,
                         padding_value=num_timestamp + 1e2)
    end = pad_sequence([torch.FloatTensor(i) for i in end_list], batch_first=True, padding_value=-1e2)
    if start.shape[1] != num_text:
        # the reference broadcasts a [B, max_len] tensor against num_text and fails when they differ
        raise RuntimeError(f"get_mask_from_time: longest start list has {start.shape[1]} entries but "
                           f"num_text={num_text} (reference train/loss.py:32-42 requires them equal)")
    steps = torch.arange(num_timestamp, dtype=torch.float32)[None, None, :]
    mask = (start[:, :, None] <= steps) & (steps < end[:, :, None])
    return (mask.to(device, non_blocking=True), start.to(device, non_blocking=True),
            end.to(device, non_blocking=True))


def get_text_pos(start_list, end_list, device='cuda'):
    """[B, N, 2] padded (start, end) (reference train/loss.py:46-54)."""
    start = pad_sequence([torch.FloatTensor(i) for i in start_list], batch_first=True, padding_value=0)
    end = pad_sequence([torch.FloatTensor(i) for i in end_list], batch_first=True, padding_value=0)
    return torch.stack((start, end), dim=-1).to(device, non_blocking=True)


def _stage_major(lg):
    """[B,S,T,Bc,N] (any strides) -> contiguous [S, B*T, Bc*N]; zero-copy for exoground_b200 outputs."""
    B, S, T, Bc, N = lg.shape
    x = lg.permute(1, 0, 2, 3, 4)
    if not x.is_contiguous():
        x = x.contiguous()
    return x.view(S, B * T, Bc * N)


def _weights(row_mask, col_mask, S, grp=None):
    """wr/wc such that sum_s,r wr*v + sum_s,c wc*t == (v[:, rows].mean() + t[:, cols].mean()) / 2.
    Cross-GPU negatives: row_mask is this rank's rows (their count is summed over ranks), col_mask is global."""
    rm, cm = row_mask.float(), col_mask.float()
    n_rows = rm.sum()
    if grp is not None:
        n_rows = grp.sum_(n_rows)
    return rm / (2.0 * S * n_rows), cm / (2.0 * S * cm.sum())


def _masked_quantile(x, mask, q):
    """torch.quantile(x[mask], q) (linear interpolation) with static shapes and no host round trip, so the
    loss stays CUDA-graph capturable: sort with +inf in the masked slots, then gather the two neighbours."""
    n = mask.sum()
    srt = torch.sort(torch.where(mask, x, torch.full_like(x, float('inf')))).values
    pos = (n - 1).to(x.dtype) * q
    lo = pos.floor().long().clamp(min=0)
    hi = torch.minimum(lo + 1, (n - 1).clamp(min=0))
    w = pos - lo.to(x.dtype)
    a, b = srt.gather(0, lo.reshape(1))[0], srt.gather(0, hi.reshape(1))[0]     # no 0-d tensor indexing (host sync)
    return a + (b - a) * w


def _masked_standardise(x, mask):
    n = mask.float().sum()
    mean = torch.where(mask, x, torch.zeros_like(x)).sum() / n
    var = torch.where(mask, (x - mean) ** 2, torch.zeros_like(x)).sum() / (n - 1)      # unbiased, as Tensor.std
    return (x - mean) / var.sqrt()


class TanLossTailFn(torch.autograd.Function):
    """loss = ((dual + joint) / 2) * nce_weight + BCE from the exp-sums of exg_nce_fwd (train/loss.py:251-376):
    forward = exg_nce_logdiff + exg_tan_loss_tail, backward = one scale of the packed weights + two exg_nce_bwd."""

    @staticmethod
    def forward(ctx, ld, lj, la, meta):
        K = ops.get()
        (tgt3, col_valid, sums_d, sums_j, md, mj, abs_c, dims, th, use_head, nce_weight) = meta
        B, T, N = dims[:3]
        out, w = K.tan_loss_tail(tgt3[0], col_valid, sums_d, sums_j, md, mj,
                                 la.contiguous() if la is not None else None, abs_c, B, T, N, INV_TEMP, th,
                                 use_head, nce_weight, tgt_bits=tgt3[1])
        ctx.saved = (ld, lj, tgt3, col_valid, sums_d, sums_j, w, dims, la is not None)
        ctx.mark_non_differentiable(out)
        return out[0].clone(), out

    @staticmethod
    def backward(ctx, g, _):
        K = ops.get()
        ld, lj, (tgt, bits, orm), col_valid, sums_d, sums_j, w, dims, has_la = ctx.saved
        B, T, N = dims[:3]
        R, Cc = B * T, B * N
        wg = w * g
        Gd = K.nce_bwd(ld, tgt, col_valid, sums_d, wg[0, :R], wg[1, :Cc], B, T, N, INV_TEMP, tgt_bits=bits)
        Gj = K.nce_bwd(lj, tgt, col_valid, sums_j, wg[2, :R], wg[3, :Cc], B, T, N, INV_TEMP, tgt_bits=bits)
        return Gd, Gj, (wg[4, :Cc] if has_la else None), None


def get_loss(input_data, video_seq, text_embed, video_padding_mask, text_padding_mask, logits, args,
             abs_text_pos=None):
    """Same contract as reference train/loss.py::get_loss for args.model == 'init'
    (returns the loss dict; 'loss' is differentiable)."""
    if args.model not in ('init', 'cotrain'):
        raise NotImplementedError("get_loss: args.model must be 'init' or 'cotrain'")
    if args.sim != 'cos':
        raise NotImplementedError("only args.sim == 'cos' is supported")
    device = logits['logits_dual'].device
    B, T, _ = video_seq.shape
    N = text_embed.shape[1]
    if '_exg_tgt' in input_data:        # device-resident target mask prepared by the caller (train/step.py)
        tgt = input_data['_exg_tgt']
    else:
        tgt, _, _ = get_mask_from_time(input_data['start'], input_data['end'], num_timestamp=T, num_text=N,
                                       device=device)                               # bool [B,N,T]
    return loss_core(tgt, text_padding_mask, logits, args, abs_text_pos, B, T, N, video_padding_mask)


def loss_core(tgt, text_padding_mask, logits, args, abs_text_pos, B, T, N, video_padding_mask=None, xneg=None):
    """Device part of get_loss: everything after the host-built target mask (static shapes, no host sync).
    xneg (train.xneg.Group, optional): cross-GPU negatives — `logits` are this rank's row blocks [B,S,T,Bg,N] of the
    global logits; the returned 'loss' is this rank's share (own rows + own columns) of the GLOBAL loss, so the sum
    over ranks equals the single-process loss on the concatenated batch; the logged entries are already summed."""
    K = ops.get()
    grp = xneg if xneg is not None and xneg.world > 1 else None
    ld, lj = _stage_major(logits['logits_dual']), _stage_major(logits['logits_joint'])
    E, D = ld.shape[0], lj.shape[0]
    Bc = ld.shape[2] // N
    clip0 = grp.rank * B if grp is not None else 0
    if Bc != (B * grp.world if grp is not None else B):
        raise ValueError(f"logits cover {Bc} column clips, expected {B} x world size")
    gather = (lambda x: grp.gather0(x)) if grp is not None else (lambda x: x)
    shard = dict(Bc=Bc, clip0=clip0)
    loss_dict = {}
    valid = ~text_padding_mask.bool()                                               # [B,N]
    valid_g = gather(valid)                                                         # [Bc,N]
    own_row_mask = None
    if getattr(args, 'learn_agreement', 0):
        # self-labelling (loss.py:91-232): targets re-derived from the (EMA) model's own last-stage logits
        from . import agreement
        if video_padding_mask is None:
            raise ValueError("learn_agreement needs video_padding_mask")
        with torch.no_grad():
            if args.model == 'cotrain':
                src_j, src_d = _stage_major(logits['ema-logits_joint']), _stage_major(logits['ema-logits_dual'])
            else:
                src_j, src_d = lj, ld
                # the reference fills the padded seconds / sentences of the own-clip blocks with -6e4 IN PLACE in
                # the logits it then trains on (loss.py:100-106,157-159); the NCE kernels reproduce that
                own_row_mask = video_padding_mask.reshape(-1).to(torch.uint8).contiguous()
            oj = agreement.own_block(src_j.detach(), src_j.shape[0] - 1, B, T, N, clip0) * INV_TEMP
            od = agreement.own_block(src_d.detach(), src_d.shape[0] - 1, B, T, N, clip0) * INV_TEMP
            tgt, stats = agreement.self_label_targets(oj, od, tgt, video_padding_mask, text_padding_mask,
                                                      args.temporal_agreement_type)
        loss_dict.update(stats)
    tgt = tgt & valid[:, :, None]
    tgt_u8 = tgt.to(torch.uint8).contiguous()
    col_valid = valid_g.reshape(-1).to(torch.uint8).contiguous()
    tgt_bits = K.nce_pack_targets(tgt_u8)          # per-(clip, second) bit rows, shared by the 4 NCE launches
    sums_d = K.nce_fwd(ld.detach(), tgt_u8, col_valid, B, T, N, INV_TEMP, tgt_bits=tgt_bits,
                       own_row_mask=own_row_mask, **shard)
    sums_j = K.nce_fwd(lj.detach(), tgt_u8, col_valid, B, T, N, INV_TEMP, tgt_bits=tgt_bits,
                       own_row_mask=own_row_mask, **shard)
    if grp is not None:                            # complete the column exp-sums over the ranks' row blocks
        for t in (sums_d[2], sums_d[3], sums_j[2], sums_j[3]):
            grp.sum_(t)
    tgt_u8 = (tgt_u8, tgt_bits, own_row_mask)
    use_th = args.loss_threshold > 0
    use_head = bool(args.use_alignability_head)
    dims = (B, T, N, Bc, clip0)

    if (use_th or use_head) and grp is None and own_row_mask is None:
        # fused tail: exp-sums -> masked means, quantile threshold, median-split labels, BCE and the backward
        # weights in two launches (csrc/loss_tail.cu) instead of ~300 elementwise / reduction / sort launches
        md = K.own_max(ld.detach(), E - 1, B, T, N)                                   # loss.py:283
        mj = K.own_max(lj.detach(), D - 1, B, T, N)                                   # loss.py:285
        la_j = abs_c = None
        if use_head:
            la_j = logits['joint_logits_alignability'][:, 2, :, 0].reshape(-1).float()   # 3rd stage hard-coded (:344)
            if abs_text_pos is not None:
                abs_c = abs_text_pos.reshape(B, N, -1).float().mean(-1).reshape(-1).contiguous()
        nce_weight = 0 if args.optim_policy == 'bce' else 1
        meta = (tgt_u8, col_valid, sums_d, sums_j, md, mj, abs_c, dims, float(args.loss_threshold) if use_th else 0.0,
                use_head, nce_weight)
        loss, out = TanLossTailFn.apply(ld, lj, la_j, meta)
        loss_dict['loss-dual'], loss_dict['loss-joint'] = out[1], out[2]
        if use_th:
            loss_dict['loss-dual-all'], loss_dict['loss-joint-all'], loss_dict['loss-total'] = out[3], out[4], out[5]
        if use_head:
            loss_dict['loss-joint-bce'], loss_dict['alignability_top1'] = out[6], out[7]
        loss_dict['loss'] = loss
        return loss_dict

    rowp = tgt.any(dim=1).reshape(-1)                                               # [B*T]   local rows
    colp = gather(tgt.any(dim=2)).reshape(-1)                                       # [Bc*N]  global columns
    # this rank's own columns: the only column terms it counts in its share of the loss value
    cidx = torch.arange(Bc * N, device=ld.device)
    own_cols = ((cidx >= clip0 * N) & (cidx < (clip0 + B) * N)).float() if grp is not None else None
    wr_d, wc_d = _weights(rowp, colp, E, grp)
    wr_j, wc_j = _weights(rowp, colp, D, grp)

    def nce(lg, sums, wr, wc):
        return NceLossFn.apply(lg, sums, tgt_u8, col_valid, wr, wc, dims, INV_TEMP, own_cols)

    if not use_th:
        loss_dual = nce(ld, sums_d, wr_d, wc_d)
        loss_joint = nce(lj, sums_j, wr_j, wc_j)
    else:
        with torch.no_grad():
            mk = dict(masked=own_row_mask is not None, n_rows=Bc * T)
            ow = (lambda w: w * own_cols) if grp is not None else (lambda w: w)
            loss_dual = nce_value(sums_d, wr_d, ow(wc_d), col_valid, INV_TEMP, **mk)
            loss_joint = nce_value(sums_j, wr_j, ow(wc_j), col_valid, INV_TEMP, **mk)
    loss_dict['loss-dual'] = loss_dual.detach()
    loss_dict['loss-joint'] = loss_joint.detach()

    if use_th or use_head:
        with torch.no_grad():
            vflat = valid_g.reshape(-1)                                              # global columns
            # the reference indexes per-text vectors compacted over *valid* columns with masks compacted over
            # columns *with positives* (loss.py:298,304,337-347): only consistent when they coincide.
            md = gather(K.own_max(ld.detach(), E - 1, B, T, N, **shard)).reshape(-1) * INV_TEMP    # loss.py:283
            mj = gather(K.own_max(lj.detach(), D - 1, B, T, N, **shard)).reshape(-1) * INV_TEMP    # loss.py:285
            metric = -(_masked_standardise(md, vflat) + _masked_standardise(mj, vflat))
            keep = vflat & (metric <= _masked_quantile(metric, vflat, args.loss_threshold))
            keep_loc = keep.view(Bc, N)[clip0:clip0 + B]
            rowp_th = (tgt & keep_loc.reshape(B, N, 1)).any(dim=1).reshape(-1)
        if use_th:
            loss_dict['loss-dual-all'] = loss_dual.detach()
            loss_dict['loss-joint-all'] = loss_joint.detach()
            wr_dt, wc_dt = _weights(rowp_th, keep, E, grp)
            wr_jt, wc_jt = _weights(rowp_th, keep, D, grp)
            loss_dual_th = nce(ld, sums_d, wr_dt, wc_dt)
            loss_joint_th = nce(lj, sums_j, wr_jt, wc_jt)
            loss_dict['loss-dual'] = loss_dual_th.detach()
            loss_dict['loss-joint'] = loss_joint_th.detach()
        if use_head:
            with torch.no_grad():
                qd, qj = _masked_quantile(md, vflat, 0.5), _masked_quantile(mj, vflat, 0.5)
                lab = torch.full_like(metric, 2.0)
                lab = torch.where((md > qd) & (mj > qj), torch.ones_like(lab), lab)
                lab = torch.where((md < qd) & (mj < qj), torch.zeros_like(lab), lab)
                if abs_text_pos is not None:
                    c = gather(abs_text_pos.reshape(B, N, -1).float().mean(-1)).reshape(-1)
                    lab = torch.where((c < 0.2) | (c > 0.8), torch.zeros_like(lab), lab)
                sel = vflat & (lab != 2)
                y = torch.where(sel, lab, torch.zeros_like(lab))
                n_sel = sel.float().sum()
                pos_weight = n_sel / y.sum() - 1.0                                    # 1/mean(y) - 1
                selw = sel.float().view(Bc, N)[clip0:clip0 + B].reshape(-1)           # this rank's sentences
                y_loc = y.view(Bc, N)[clip0:clip0 + B].reshape(-1)
            la_j = logits['joint_logits_alignability'][:, 2, :, 0].reshape(-1).float()   # 3rd stage hard-coded (:344)
            bce = F.binary_cross_entropy_with_logits(la_j, y_loc, pos_weight=pos_weight, reduction='none')
            loss_bce_joint = (bce * selw).sum() / n_sel
            loss_dict['loss-joint-bce'] = loss_bce_joint.detach()
            loss_dict['alignability_top1'] = ((((la_j > 0).float() == y_loc).float() * selw).sum() / n_sel).detach()

    nce_weight = 0 if args.optim_policy == 'bce' else 1
    if use_th:
        loss_dict['loss-total'] = ((loss_dual + loss_joint) / 2).detach()
        loss = (loss_dual_th + loss_joint_th) / 2
    else:
        loss = (loss_dual + loss_joint) / 2
    if use_head:
        loss = loss * nce_weight + loss_bce_joint
    if grp is not None:
        # logged entries: per-rank shares -> global values, one small all-reduce
        keys = [k for k in loss_dict if k not in ('confidence-ratio', 'iou-threshold')]
        if keys:
            packed = grp.sum_(torch.stack([loss_dict[k].float() for k in keys]))
            for i, k in enumerate(keys):
                loss_dict[k] = packed[i]
        loss_dict['loss-global'] = grp.sum_(loss.detach().clone())
    loss_dict['loss'] = loss
    return loss_dict
