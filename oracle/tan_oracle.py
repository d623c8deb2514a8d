"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Functional fp32 restatement of TemporalAligner (model/tan_model.py).  The
random positional offsets the reference draws with ``np.random.randint`` inside
forward (tan_model.py:157,189,218) are explicit arguments here; use
``draw_pos_starts`` to reproduce the reference's draw order from a seeded
``np.random`` state.
"""
import numpy as np
import torch

from .tfm_oracle import layer_norm, temporal_encoder, interp_pos


def draw_pos_starts(T, N, random_pos_start=1, use_text_pos_enc=0, interpolate_from=None):
    """np.random draws of one TemporalAligner.forward, in reference order:
    get_visual_feature (:157) -> [get_textual_feature_with_time (:218)] -> get_joint_feature (:189).
    Returns (p_video, p_text, p_joint)."""
    if interpolate_from or not random_pos_start:
        return 0, 0, 0
    pv = int(np.random.randint(0, int(T / 2)))
    pt = int(np.random.randint(0, int(N / 2))) if use_text_pos_enc else 0
    pj = int(np.random.randint(0, int(T / 2)))
    return pv, pt, pj


def _pos_rows(table, p0, size, interpolate_from):
    if interpolate_from:
        return interp_pos(table, interpolate_from, size)
    return table[p0:p0 + size]


def video_embed_with_time(sd, video, p0, interpolate_from=None):
    """ln_video_init(video_pre_proj(v)) + ln_position_init(pos[p0:p0+T]) (tan_model.py:149-161)."""
    x = layer_norm(video @ sd["video_pre_proj.weight"].T, sd["ln_video_init.weight"], sd["ln_video_init.bias"])
    pos = _pos_rows(sd["temporal_pos_embed"], p0, video.shape[1], interpolate_from)
    return x + layer_norm(pos, sd["ln_position_init.weight"], sd["ln_position_init.bias"])[None]


def textual_feature(sd, text):
    """get_textual_feature (tan_model.py:225-228)."""
    return layer_norm(text @ sd["text_pre_proj.weight"].T, sd["ln_text_init.weight"], sd["ln_text_init.bias"])


def textual_feature_with_time(sd, text, p0, interpolate_from=None):
    """get_textual_feature_with_time (tan_model.py:206-222)."""
    x = textual_feature(sd, text)
    pos = _pos_rows(sd["text_temporal_pos_embed"], p0, text.shape[1], interpolate_from)
    return x + layer_norm(pos, sd["ln_position_init.weight"], sd["ln_position_init.bias"])[None]


def visual_feature(sd, layers, video, vmask, p0, interpolate_from=None):
    """get_visual_feature (tan_model.py:146-173) -> [B,E,T,C]."""
    x = video_embed_with_time(sd, video, p0, interpolate_from)
    st = temporal_encoder(x, sd, "video_temporal_encoder.", layers, vmask)
    st[-1] = layer_norm(st[-1], sd["ln_video_post_enc.weight"], sd["ln_video_post_enc.bias"])
    return torch.stack(st, dim=1)


def joint_feature(sd, layers, video, vmask, text_feat, tmask, p0, interpolate_from=None):
    """get_joint_feature (tan_model.py:176-203) -> ([B,D,T,C], [B,D,N,C])."""
    T = video.shape[1]
    x = torch.cat((video_embed_with_time(sd, video, p0, interpolate_from), text_feat), dim=1)
    m = torch.cat((vmask, tmask), dim=1)
    st = temporal_encoder(x, sd, "joint_temporal_encoder.", layers, m)
    st[-1] = layer_norm(st[-1], sd["ln_joint_post_enc.weight"], sd["ln_joint_post_enc.bias"])
    st = torch.stack(st, dim=1)
    return st[:, :, :T], st[:, :, T:]


def l2n(x):
    """x / x.norm(dim=-1) with NO eps (tan_model.py:110-111)."""
    return x / x.norm(dim=-1, keepdim=True)


def tan_forward(sd, num_encoder_layers, num_decoder_layers, video, text, vmask, tmask,
                pos_starts=(0, 0, 0), use_text_pos_enc=0, use_alignability_head=0,
                interpolate_from=None):
    """TemporalAligner.forward (tan_model.py:94-143).  Returns the same dict."""
    pv, pt, pj = pos_starts
    v = visual_feature(sd, num_encoder_layers, video, vmask, pv, interpolate_from)
    t_raw = textual_feature(sd, text)
    vn, tn = l2n(v), l2n(t_raw)
    logits_dual = torch.einsum("astc,bkc->astbk", vn, tn)
    t_time = textual_feature_with_time(sd, text, pt, interpolate_from) if use_text_pos_enc else t_raw
    jv, jt = joint_feature(sd, num_decoder_layers, video, vmask, t_time, tmask, pj, interpolate_from)
    logits_joint = torch.einsum("astc,bskc->astbk", l2n(jv), l2n(jt))
    out = {"logits_dual": logits_dual, "logits_joint": logits_joint,
           "dual_feature_video": vn, "dual_feature_text": tn}
    if use_alignability_head:
        w, b = sd["binary_head.weight"], sd["binary_head.bias"]
        out["dual_logits_alignability"] = t_raw @ w.T + b
        out["joint_logits_alignability"] = jt @ w.T + b
    return out


def text_visual_sim_joint(sd, layers, video, text, p0=0, interpolate_from=None):
    """get_text_visual_sim_joint without text pos-enc (tan_model.py:231-258) -> [B,S,T,K]."""
    B, T, _ = video.shape
    N = text.shape[1]
    vm = torch.zeros(B, T, dtype=torch.bool)
    tm = torch.zeros(B, N, dtype=torch.bool)
    jv, jt = joint_feature(sd, layers, video, vm, textual_feature(sd, text), tm, p0, interpolate_from)
    return torch.einsum("bstc,bskc->bstk", l2n(jv), l2n(jt))


def text_visual_sim_dual(sd, layers, video, text, p0=0, interpolate_from=None):
    """get_text_visual_sim_dual (tan_model.py:261-277) -> [B,S,T,K]."""
    B, T, _ = video.shape
    vm = torch.zeros(B, T, dtype=torch.bool)
    v = visual_feature(sd, layers, video, vm, p0, interpolate_from)
    return torch.einsum("bstc,bkc->bstk", l2n(v), l2n(textual_feature(sd, text)))
