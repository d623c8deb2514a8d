"""Drop-in replacement of the reference's model/keystep_ground_model.py::GroundingModel: the ExoGround transformer
whose video input first passes through a frozen, pre-trained view-invariant encoder (`vi_encoder`, run under
no_grad; keystep_ground_model.py:158-169) and whose distillation InfoNCE is evaluated inside forward() and
returned as `distill_infonce_loss` (:174-176,200-201).  Same constructor kwargs, parameter names (the frozen
encoder's parameters appear under `vi_encoder.`), forward signature and output dict.

The reference's TwinExoGroundingTransformer (:411-448) references attributes that do not exist on the model
(`online.bert`, `get_visual_feature`, ...) and cannot be constructed; it is not reproduced.
"""
import torch

from . import functional as Fn
from .exo_ground_model import ExoGroundingTransformer, info_nce_bmm, pairwise_info_nce


class GroundingModel(ExoGroundingTransformer):
    def __init__(self, *args, vi_encoder=None, **kwargs):
        if len(args) > 17:                       # vi_encoder is the 18th positional argument of the reference
            vi_encoder, args = args[17], args[:17]
        super().__init__(*args, **kwargs)
        self.vi_encoder = vi_encoder             # a registered sub-module: its weights stay in state_dict()

    def forward(self, video_embed, lang_embed, video_padding_mask, lang_padding_mask, audio_embed=None,
                audio_padding_mask=None, egocentric_video_embed=None, view_mask=None, interpolate_from=None):
        self._check(video_embed, lang_embed, audio_embed)
        wT = self._weights()
        B, T, _ = video_embed.shape
        N = lang_embed.shape[1]
        vmask = video_padding_mask.to(torch.uint8).contiguous()
        tmask = lang_padding_mask.to(torch.uint8).contiguous()
        xt = self._text_proj(lang_embed, wT)
        a_feat = amask = None
        if audio_embed is not None:
            a_feat = self._audio_proj(audio_embed, wT)
            amask = audio_padding_mask.to(torch.uint8).contiguous()
        if self.use_text_pos_enc:
            p0 = 0 if interpolate_from else self._draw(N)
            xt = Fn.AddPosFn.apply(xt, self._pos_ln(self.text_temporal_pos_embed, p0, N, interpolate_from), B, N)
        if self.vi_encoder is not None:
            with torch.no_grad():
                vi_video = self.vi_encoder(video_embed, lang_embed, video_padding_mask, lang_padding_mask,
                                           audio_embed=audio_embed, audio_padding_mask=audio_padding_mask,
                                           egocentric_video_embed=egocentric_video_embed, view_mask=view_mask,
                                           interpolate_from=interpolate_from)['high_dim_features']
        else:
            vi_video = video_embed
        v_enc, t_enc = self._unimodal_pair(vi_video, vmask, xt, tmask, B, N, wT)
        distill = None
        if self.use_distill_nce_loss and egocentric_video_embed is not None:
            distill = info_nce_bmm(self._exo_proj(v_enc, B, T, wT), egocentric_video_embed, self.compute_dtype)
        elif self.multi_view and self.use_pairwise_distill_nce_loss:          # reference :177-179
            distill = pairwise_info_nce(self._exo_proj(v_enc, B, T, wT),
                                        view_mask if self.pairwise_distill_mode == "all" else ~video_padding_mask,
                                        self.num_max_views)
        grounding = self._grounding(v_enc, t_enc, vmask, tmask, B, T, N, wT, a_feat, amask, interpolate_from)
        out = {'interval_preds': grounding}
        if self.use_distill_nce_loss or self.use_pairwise_distill_nce_loss:
            if distill is None:
                raise NameError("distill_loss is undefined without egocentric_video_embed "
                                "(reference keystep_ground_model.py:174-179,200-201 fails the same way)")
            out['distill_infonce_loss'] = distill
        return out
