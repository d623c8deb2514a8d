"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Functional fp32 restatement of ExoGroundingTransformer (model/exo_ground_model.py),
GroundingModel (model/keystep_ground_model.py), ViewInvariantMLP (model/vi_encoder.py:166-256)
and the ExoGround losses (train/loss_egoexo4d.py:73-363).
"""
import numpy as np
import torch
import torch.nn.functional as F

from .tfm_oracle import layer_norm, temporal_encoder, temporal_decoder, interp_pos


def draw_pos_starts(T, num_max_views=1, random_pos_start=1, interpolate_from=None):
    """np.random draws of one ExoGroundingTransformer.forward without text pos-enc:
    get_unimodal_features('video') (exo_ground_model.py:305) then get_joint_feature (:338, value unused)."""
    if interpolate_from or not random_pos_start:
        return 0, 0
    sl = T // num_max_views
    return int(np.random.randint(0, int(sl / 2))), int(np.random.randint(0, int(sl / 2)))


def textual_feature(sd, text):
    """exo_ground_model.py:385-388."""
    return layer_norm(text @ sd["text_pre_proj.weight"].T, sd["ln_text_init.weight"], sd["ln_text_init.bias"])


def unimodal_video(sd, layers, video, vmask, p0=0, num_max_views=1, interpolate_from=None):
    """get_unimodal_features('video', ...) (exo_ground_model.py:294-321) -> [B,E,T,C]."""
    T = video.shape[1]
    sl = T // num_max_views
    x = layer_norm(video @ sd["video_pre_proj.weight"].T, sd["ln_video_init.weight"], sd["ln_video_init.bias"])
    tab = sd["temporal_pos_embed"]
    pos = interp_pos(tab, interpolate_from, sl) if interpolate_from else tab[p0:p0 + sl]
    pos = pos.repeat(num_max_views, 1)
    x = x + layer_norm(pos, sd["ln_position_init.weight"], sd["ln_position_init.bias"])[None]
    st = temporal_encoder(x, sd, "video_unimodal_encoder.", layers, vmask)
    st[-1] = layer_norm(st[-1], sd["ln_video_post_enc.weight"], sd["ln_video_post_enc.bias"])
    return torch.stack(st, dim=1)


def unimodal_text(sd, layers, text_feat, tmask):
    """get_unimodal_features('text', ...) (exo_ground_model.py:314-321) -> [B,E,N,C]."""
    st = temporal_encoder(text_feat, sd, "text_unimodal_encoder.", layers, tmask)
    st[-1] = layer_norm(st[-1], sd["ln_text_post_enc.weight"], sd["ln_text_post_enc.bias"])
    return torch.stack(st, dim=1)


def exo_forward(sd, num_encoder_layers, num_decoder_layers, video, text, vmask, tmask,
                p0=0, use_decoder=True, use_distill=False, ego=None, num_max_views=1,
                interpolate_from=None, vi_sd=None):
    """ExoGroundingTransformer.forward (exo_ground_model.py:127-183) without audio / text pos-enc.
    With vi_sd given, the GroundingModel variant (keystep_ground_model.py:158-171,174-179,199-201)."""
    t_raw = textual_feature(sd, text)
    N = t_raw.shape[1]
    vin = video
    if vi_sd is not None:
        vin = vi_mlp_forward(vi_sd, video)
    v_enc = unimodal_video(sd, num_encoder_layers, vin, vmask, p0, num_max_views, interpolate_from).mean(dim=1)
    t_enc = unimodal_text(sd, num_encoder_layers, t_raw, tmask).mean(dim=1)
    x = torch.cat((v_enc, t_enc), dim=1)
    m = torch.cat((vmask, tmask), dim=1)
    st = temporal_encoder(x, sd, "multi_modal_encoder.", num_encoder_layers, m)
    st[-1] = layer_norm(st[-1], sd["ln_joint_post_enc.weight"], sd["ln_joint_post_enc.bias"])
    last = st[-1]
    ctx, txt = last[:, :-N], last[:, -N:]
    if use_decoder:
        dec = temporal_decoder(txt, ctx, sd, "decoder.", num_decoder_layers, tmask, vmask)
        g = dec[-1] @ sd["grounding_head.weight"].T + sd["grounding_head.bias"]
    else:
        # reference applies the head to ALL stages here (exo_ground_model.py:177): [B,E,N,2]
        g = torch.stack(st, 1)[:, :, -N:] @ sd["grounding_head.weight"].T + sd["grounding_head.bias"]
    out = {"interval_preds": g, "low_dim_features": v_enc}
    if use_distill:
        proj = v_enc @ sd["exo_feature_proj.weight"].T + sd["exo_feature_proj.bias"]
        if vi_sd is not None or ego is None:
            pass
        out["high_dim_features"] = proj
        if vi_sd is not None and ego is not None:
            out["distill_infonce_loss"] = info_nce_bmm(proj, ego)
    return out


def vi_mlp_forward(sd, video):
    """ViewInvariantMLP.forward (vi_encoder.py:245-256)."""
    x = layer_norm(video @ sd["video_pre_proj.weight"].T, sd["ln_video_init.weight"], sd["ln_video_init.bias"])
    x = torch.relu(x @ sd["mlp.0.weight"].T + sd["mlp.0.bias"])
    return x @ sd["mlp.2.weight"].T + sd["mlp.2.bias"]


def info_nce_bmm(f1, f2, temperature=0.1):
    """compute_info_nce_loss (exo_ground_model.py:206-234; loss_egoexo4d.py:325-347)."""
    a = F.normalize(f1, p=2, dim=2)
    b = F.normalize(f2, p=2, dim=2)
    sim = a @ b.transpose(1, 2) / temperature
    lp = torch.log_softmax(sim, dim=2)
    return -torch.diagonal(lp, dim1=1, dim2=2).mean()


def grounding_loss(preds, gt_a, gt_b, text_pad, use_center_duration=True, eps=1e-8,
                   thresholds=(0.1, 0.3, 0.5, 0.7)):
    """get_grounding_loss_reg_head (loss_egoexo4d.py:73-145), train-mode keys only."""
    valid = ~text_pad.bool()
    pa, pb = preds[:, :, 0][valid], preds[:, :, 1][valid]
    ga, gb = gt_a[valid], gt_b[valid]
    l1a, l1b = (pa - ga).abs().mean(), (pb - gb).abs().mean()
    out = {}
    if use_center_duration:
        ps, pe, gs, ge = pa - pb / 2, pa + pb / 2, ga - gb / 2, ga + gb / 2
        out["Center L1 loss"], out["Duration L1 loss"] = l1a, l1b
    else:
        ps, pe, gs, ge = pa, pb, ga, gb
        out["Timestamp L1 loss"] = (l1a + l1b) / 2
    inter = (torch.minimum(pe, ge) - torch.maximum(ps, gs)).clamp(min=0)
    union = torch.maximum(pe, ge) - torch.minimum(ps, gs)
    iou = inter / (union + eps)
    out["IoU loss"] = 1.0 - iou.mean()
    out["mean IoU"] = iou.mean()
    for th in thresholds:
        out[f"IoU>={th}"] = (iou > th).sum().item() / valid.sum().item()
    loss = out["IoU loss"].clone()
    loss = loss + (out["Duration L1 loss"] + out["Center L1 loss"] if use_center_duration
                   else out["Timestamp L1 loss"])
    out["loss"] = loss
    return out, iou


def _gather_view(feats, idx):
    """feats [B,V,T,D], idx [B,T] -> [B,T,D] where row (b,t) = feats[b, idx[b,t0?]...].

    NB the reference's gather (loss_egoexo4d.py:283-287) expands idx [B,T] as
    idx.unsqueeze(1).unsqueeze(-1).expand(-1,-1,T,D): index tensor [B,1,T,D] with value idx[b,t],
    so out[b,t,:] = feats[b, idx[b,t], t, :]."""
    B, V, T, D = feats.shape
    return feats[torch.arange(B)[:, None], idx, torch.arange(T)[None, :]]


def xview_info_nce(out_feat, ego, pos_idx, neg_idx, same_idx=None, same_feat=None,
                   only_same_view_negative=False, temperature=0.1):
    """compute_info_nce_loss_cross_view (loss_egoexo4d.py:269-323) -> per-(b,t) loss [B,T]."""
    o = F.normalize(out_feat, p=2, dim=2)
    p = F.normalize(_gather_view(ego, pos_idx), p=2, dim=2)
    n = F.normalize(_gather_view(ego, neg_idx), p=2, dim=2)
    sims = [(o * p).sum(2) / temperature, (o * n).sum(2) / temperature]
    if same_idx is not None:
        B, T, D = same_feat.shape
        s = F.normalize(same_feat[torch.arange(B)[:, None], same_idx], p=2, dim=2)
        ss = (o * s).sum(2) / temperature
        sims = [sims[0], ss] if only_same_view_negative else sims + [ss]
    return -torch.log_softmax(torch.stack(sims, 2), dim=2)[:, :, 0]


def l1_cosine(out_feat, ego, pos_idx, neg_idx):
    """compute_l1_cosine_losses (loss_egoexo4d.py:201-233)."""
    o = F.normalize(out_feat, p=2, dim=2)
    p = F.normalize(_gather_view(ego, pos_idx), p=2, dim=2)
    n = F.normalize(_gather_view(ego, neg_idx), p=2, dim=2)
    return (o - p).abs().mean(2), (o * p).sum(2), (o * n).sum(2)
