"""Parameter name -> shape tables of the reference modules (model/tan_model.py:37-67,
model/exo_ground_model.py:56-103, model/vi_encoder.py:211-221), written out so tests can build
state dicts on a box without /root/reference.  test_shapes_match_reference checks them against the
real modules when the reference tree is present."""
C, F_ = 512, 2048


def _enc(prefix, layers, out):
    for i in range(layers):
        p = f"{prefix}.resblocks.{i}."
        out[p + "attn.in_proj_weight"] = (3 * C, C)
        out[p + "attn.in_proj_bias"] = (3 * C,)
        out[p + "attn.out_proj.weight"] = (C, C)
        out[p + "attn.out_proj.bias"] = (C,)
        out[p + "ln_1.weight"] = (C,)
        out[p + "ln_1.bias"] = (C,)
        out[p + "mlp.c_fc.weight"] = (F_, C)
        out[p + "mlp.c_fc.bias"] = (F_,)
        out[p + "mlp.c_proj.weight"] = (C, F_)
        out[p + "mlp.c_proj.bias"] = (C,)
        out[p + "ln_2.weight"] = (C,)
        out[p + "ln_2.bias"] = (C,)


def _dec(prefix, layers, out):
    _enc(prefix, layers, out)
    for i in range(layers):
        p = f"{prefix}.resblocks.{i}."
        out[p + "self_attn.in_proj_weight"] = (3 * C, C)
        out[p + "self_attn.in_proj_bias"] = (3 * C,)
        out[p + "self_attn.out_proj.weight"] = (C, C)
        out[p + "self_attn.out_proj.bias"] = (C,)
        out[p + "ln_3.weight"] = (C,)
        out[p + "ln_3.bias"] = (C,)


def _ln(names, out):
    for n in names:
        out[n + ".weight"] = (C,)
        out[n + ".bias"] = (C,)


def tan_shapes(E, D, d_in=4096, head=0, pos_enc="learned"):
    out = {"temporal_pos_embed": (4096, C), "text_temporal_pos_embed": (4096, C)}
    _enc("video_temporal_encoder", E, out)
    _enc("joint_temporal_encoder", D, out)
    out["video_pre_proj.weight"] = (C, d_in)
    out["text_pre_proj.weight"] = (C, d_in)
    _ln(["ln_text_init", "ln_video_init", "ln_position_init", "ln_video_post_enc", "ln_joint_post_enc"], out)
    out["mlp.weight"] = (C, C)
    out["mlp.bias"] = (C,)
    if head:
        out["binary_head.weight"] = (1, C)
        out["binary_head.bias"] = (1,)
    return out


def exo_shapes(E, D, d_video=4096, d_text=4096, distill=False, use_decoder=True, d_audio=0):
    out = {"temporal_pos_embed": (1024, C), "text_temporal_pos_embed": (d_text, C)}
    _enc("multi_modal_encoder", E, out)
    if use_decoder:
        _dec("decoder", D, out)
    out["grounding_head.weight"] = (2, C)
    out["grounding_head.bias"] = (2,)
    _enc("video_unimodal_encoder", E, out)
    _enc("text_unimodal_encoder", E, out)
    out["video_pre_proj.weight"] = (C, d_video)
    out["text_pre_proj.weight"] = (C, d_text)
    _ln(["ln_text_init", "ln_video_init", "ln_position_init", "ln_joint_post_enc",
         "ln_video_post_enc", "ln_text_post_enc"], out)
    if distill:
        out["exo_feature_proj.weight"] = (d_video, C)
        out["exo_feature_proj.bias"] = (d_video,)
    if d_audio:
        _ln(["ln_audio_init"], out)
        out["audio_pre_proj.weight"] = (C, d_audio)
    out["mlp.weight"] = (C, C)
    out["mlp.bias"] = (C,)
    return out


def vi_shapes(d_video=4096):
    D = d_video
    return {"video_pre_proj.weight": (D, D), "ln_video_init.weight": (D,), "ln_video_init.bias": (D,),
            "mlp.0.weight": (D, D), "mlp.0.bias": (D,), "mlp.2.weight": (D, D), "mlp.2.bias": (D,)}
