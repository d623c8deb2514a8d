"""Drop-in replacement of the reference's train/loss_egoexo4d.py loss entry points (get_loss and the two loss
families it dispatches to: loss_egoexo4d.py:73-145 interval regression + IoU, :150-199,269-323 cross-view InfoNCE
with L1 / cosine monitors, :325-347 per-clip InfoNCE).  Same signatures, same loss_dict keys.  The arithmetic
runs in libexg.so (exg_ground_loss: one launch for the masked L1 + IoU + threshold counts and their gradient;
exg_xview_nce_fwd/bwd: one warp per (clip, second)); the host code below only does the reference's bookkeeping
(camera-rank modes, per-rank bins).  The reference's visualize() (matplotlib / ffmpeg / cv2, :365-) is out of scope.
"""
from collections import Counter

import numpy as np
import torch
from torch.nn.utils.rnn import pad_sequence

from ..model import functional as Fn
from ..model.exo_ground_model import info_nce_bmm
from .loss import circulant  # noqa: F401  (same helper, loss_egoexo4d.py:18-26)

INV_TEMP = 1.0 / 0.1


def get_mask_from_time(start_list, end_list, num_timestamp, num_text, device='cuda'):
    """[B, num_text, T] mask of seconds covered by each sentence (loss_egoexo4d.py:29-46)."""
    B = len(start_list)
    steps = torch.arange(num_timestamp, device=device)[None, None, :].repeat(B, num_text, 1)
    start = pad_sequence([torch.FloatTensor(i) for i in start_list], batch_first=True,
                         padding_value=num_timestamp + 1e2).to(device, non_blocking=True)
    end = pad_sequence([torch.FloatTensor(i) for i in end_list], batch_first=True,
                       padding_value=-1e2).to(device, non_blocking=True)
    mask = (start[:, :, None] <= steps) * (steps < end[:, :, None])
    return mask, start, end


def get_text_pos(start_list, end_list, device='cuda'):
    """loss_egoexo4d.py:49-57."""
    start = pad_sequence([torch.FloatTensor(i) for i in start_list], batch_first=True, padding_value=0)
    end = pad_sequence([torch.FloatTensor(i) for i in end_list], batch_first=True, padding_value=0)
    return torch.stack((start.to(device, non_blocking=True), end.to(device, non_blocking=True)), dim=-1)


def get_mode_cam_rank(batch_cam_rank_metadata):
    """Most frequent camera rank of each clip; ties go to the first seen (loss_egoexo4d.py:59-65)."""
    modes = []
    for ranks in batch_cam_rank_metadata:
        count = Counter(ranks)
        modes.append(max(count, key=count.get))
    return modes


def expand_ranks_with_mask(modes, text_padding_mask):
    """One entry per un-padded sentence, clip-major (loss_egoexo4d.py:67-75)."""
    n_valid = (~text_padding_mask.bool()).sum(dim=1).tolist()
    out = []
    for mode, n in zip(modes, n_valid):
        out.extend([mode] * int(n))
    return out


def flatten_list_of_lists(list_of_lists):
    return [item for sub in list_of_lists for item in sub]


def ground_loss_core(preds, gt_a, gt_b, text_padding_mask, use_center_duration, eps, thresholds):
    """Device part of get_grounding_loss_reg_head: preds [B,N,2] (autograd), gt [B,N], mask [B,N].
    Returns (loss, stats, iou_all): stats = (mean|d0|, mean|d1|, mean IoU, n_valid, count(IoU > th)...),
    iou_all [B*N] (0 on padded sentences).  No host synchronisation (CUDA-graph safe)."""
    B, N, _ = preds.shape
    return Fn.GroundLossFn.apply(preds.reshape(B * N, 2), gt_a.reshape(-1).float().contiguous(),
                                 gt_b.reshape(-1).float().contiguous(),
                                 text_padding_mask.reshape(-1).to(torch.uint8).contiguous(),
                                 bool(use_center_duration), float(eps), tuple(float(t) for t in thresholds))


def grounding_dict(stats, host, use_center_duration, thresholds):
    """The logged entries of get_grounding_loss_reg_head (loss_egoexo4d.py:102-137) from the kernel's statistics:
    `stats` = tensor (mean|d0|, mean|d1|, mean IoU, n_valid, counts...) on any device, `host` = the same as floats."""
    d = {}
    if use_center_duration:
        d['Center L1 loss'] = stats[0]
        d['Duration L1 loss'] = stats[1]
    else:
        d['Timestamp L1 loss'] = (stats[0] + stats[1]) / 2
    d['IoU loss'] = 1.0 - stats[2]
    d['mean IoU'] = stats[2]
    n_valid = host[3]
    for i, theta in enumerate(thresholds):
        d[f'IoU>={theta}'] = host[4 + i] / n_valid
    return d


def view_invariant_dict(host, view_rank_names, means):
    """The logged entries of get_view_invariant_loss (loss_egoexo4d.py:173-198): per-camera-rank bins from the
    per-(clip, second) values `host` = float64 [4, B*T] (L1, pos cosine, neg cosine, InfoNCE) and the four means."""
    d = {}
    names, inverse = np.unique(np.asarray(flatten_list_of_lists(view_rank_names), dtype=object).astype(str),
                               return_inverse=True)
    counts = np.bincount(inverse, minlength=len(names))
    for metric, vals in zip(("L1", "pos_cosine", "avg_neg_cosine", "InfoNCE"), host):
        totals = np.bincount(inverse, weights=vals, minlength=len(names))        # float64 sums, as the reference's
        for name, tot, cnt in zip(names, totals, counts):
            d[f"Rank {name} {metric}"] = {'total': float(tot), 'count': int(cnt), 'mean': float(tot / cnt)}
    d["L1 loss"], d["Pos cosine sim"], d["Avg neg cosine sim"], d['InfoNCE loss'] = means
    return d


def get_grounding_loss_reg_head(input_data, logits, text_padding_mask, args):
    """loss_egoexo4d.py:73-145."""
    preds = logits['interval_preds']
    device = preds.device
    text_padding_mask = text_padding_mask.to(device)
    if args.use_center_duration:
        gt_a = input_data['mean'].to(device, non_blocking=True)
        gt_b = input_data['duration'].to(device, non_blocking=True)
    else:
        gt_a = input_data['starts'].to(device, non_blocking=True)
        gt_b = input_data['ends'].to(device, non_blocking=True)
    thresholds = list(args.iou_thresholds)
    loss, stats, iou_all = ground_loss_core(preds, gt_a, gt_b, text_padding_mask, args.use_center_duration,
                                            args.iou_loss_eps, thresholds)
    loss_dict = grounding_dict(stats, stats.tolist(), args.use_center_duration, thresholds)   # one read-back (:121-122)
    iou = iou_all[~text_padding_mask.reshape(-1).bool()]
    if getattr(args, 'test', False):
        cam_ranks = get_mode_cam_rank(input_data['metadata']['per_second_views'])
        expanded = expand_ranks_with_mask(cam_ranks, text_padding_mask.cpu())
        iou_host = iou.detach().cpu()
        for cam_rank in set(expanded):
            rank_iou = iou_host[torch.tensor([r == cam_rank for r in expanded])]
            for theta in thresholds:
                loss_dict[f'Rank {cam_rank} IoU>={theta}'] = {
                    'mean': (rank_iou > theta).sum().item() / len(rank_iou), 'count': len(rank_iou)}
    loss_dict['loss'] = loss
    return loss_dict, iou


def compute_info_nce_loss_cross_view(output_features, video_features, positive_indices, negative_indices,
                                     same_view_neg_idxs=None, same_view_features=None,
                                     only_same_view_negative=False, temperature=0.1):
    """Per-(clip, second) InfoNCE of the output feature against the best-view (positive) and a worse-view
    (negative) ego-exo feature of the same second, optionally a same-view negative from another second
    (loss_egoexo4d.py:269-323) -> [B, T]."""
    return _xview(output_features, video_features, positive_indices, negative_indices, same_view_neg_idxs,
                  same_view_features, only_same_view_negative, 1.0 / temperature)[0]


def compute_l1_cosine_losses(output_features, video_features, positive_indices, negative_indices):
    """(L1 between normalised output and positive, cos to positive, cos to negative), each [B,T]
    (loss_egoexo4d.py:201-233); monitors only."""
    with torch.no_grad():
        return _xview(output_features, video_features, positive_indices, negative_indices, None, None, False,
                      INV_TEMP)[1:]


def _xview(out_feat, ego, pos_idx, neg_idx, same_idx, same_feat, only_same, inv_temp):
    ego = ego.float().contiguous()
    if same_idx is not None:
        same_feat, same_idx = same_feat.float().contiguous(), same_idx.long().contiguous()
    else:
        same_feat = None
    return Fn.XviewNceFn.apply(out_feat.float().contiguous(), ego, pos_idx.long().contiguous(),
                               neg_idx.long().contiguous(), same_feat, same_idx, bool(only_same), float(inv_temp))


def compute_info_nce_loss(features1, features2, temperature=0.1):
    """loss_egoexo4d.py:325-347."""
    cd = torch.bfloat16 if features1.dtype == torch.bfloat16 else torch.float32
    return info_nce_bmm(features1, features2, cd, 1.0 / temperature)


def get_view_invariant_loss(input_data, logits, args):
    """loss_egoexo4d.py:150-199."""
    features = logits['high_dim_features'] if not getattr(args, 'test_egovlp', False) else input_data['video_features']
    device = features.device
    view_rank_names = input_data['metadata']['per_second_views']
    ego_seq = input_data['ego_video_features'].to(device, non_blocking=True)
    pos_idx = input_data['view_rank_label'].to(device, non_blocking=True)
    neg_idx = input_data['view_rank_neg_label'].to(device, non_blocking=True)
    if getattr(args, 'same_view_negative', False):
        same_idx = input_data['same_view_neg_idxs'].to(device, non_blocking=True)
        same_feat = input_data['video_features'].to(device, non_blocking=True)
    else:
        same_idx = same_feat = None
    nce, l1, pos_cos, neg_cos = _xview(features, ego_seq, pos_idx, neg_idx, same_idx, same_feat,
                                       getattr(args, 'only_same_view_negative', False), INV_TEMP)
    # per-camera-rank bins (reference :173-187, one .item() per element there; one read-back here)
    host = torch.stack((l1, pos_cos, neg_cos, nce.detach()), dim=0).reshape(4, -1).double().cpu().numpy()
    if not args.use_distill_nce_loss:
        raise UnboundLocalError("total_loss is undefined unless args.use_distill_nce_loss "
                                "(reference loss_egoexo4d.py:195-198 fails the same way)")
    loss_dict = view_invariant_dict(host, view_rank_names, (l1.mean(), pos_cos.mean(), neg_cos.mean(), nce.mean()))
    loss_dict['loss'] = loss_dict['InfoNCE loss'].clone()
    return loss_dict, None


def get_loss(input_data, logits, text_padding_mask, args):
    """loss_egoexo4d.py:349-363."""
    if args.model in ['view_invariant']:
        return get_view_invariant_loss(input_data, logits, args)
    if args.model in ['grounding'] or (args.model in ['joint'] and not args.use_distill_nce_loss):
        return get_grounding_loss_reg_head(input_data, logits, text_padding_mask, args)
    if args.model in ['joint']:
        gnd, iou = get_grounding_loss_reg_head(input_data, logits, text_padding_mask, args)
        vi, _ = get_view_invariant_loss(input_data, logits, args)
        combined = {**vi, **gnd}
        combined['loss'] = vi['loss'] + gnd['loss']
        return combined, iou
    return None
