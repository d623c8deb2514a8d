"""Drop-in replacement of the reference's model/vi_encoder.py: ViewInvariantMLP (4096 -> 4096 projection, LayerNorm,
Linear-ReLU-Linear; vi_encoder.py:171-253) and ViewInvariantEncoder (video unimodal encoder + stage mean +
512 -> D_v projection; vi_encoder.py:8-168).  Same constructor kwargs, parameter names, forward signature and
output dict.  Used as the `view_invariant` model of train/main_egoexo4d_distributed.py and, frozen, as the
`vi_encoder` of keystep_ground_model.GroundingModel.
"""
import numpy as np
import torch
from torch import nn
from torch.nn import LayerNorm

from .. import ops
from . import functional as Fn
from .params import ExgModule
from .tfm_model import TemporalEncoder, get_position_embedding_sine, interp_rows

_KW = dict(num_encoder_layers=2, num_decoder_layers=2, use_decoder=True, sim='cos', pos_enc='learned',
           use_text_pos_enc=0, random_pos_start=1, use_audio=False, video_embed_dim=4096, text_embed_dim=4096,
           audio_embed_dim=2304, feature_dim=512, use_distill_nce_loss=False, multi_view=False, num_max_views=1,
           use_pairwise_distill_nce_loss=False, pairwise_distill_mode="all")


def _take_kwargs(self, args, kwargs):
    """The reference constructors share one 17-keyword signature (vi_encoder.py:9-27,172-190); store them all."""
    names = list(_KW)
    if len(args) > len(names):
        raise TypeError("too many positional arguments")
    cfg = dict(_KW)
    cfg.update(dict(zip(names, args)))
    for k, v in kwargs.items():
        if k not in cfg and k != "compute_dtype":
            raise TypeError(f"unexpected keyword argument {k!r}")
        if k in cfg:
            cfg[k] = v
    for k, v in cfg.items():
        setattr(self, k, v)
    return kwargs.get("compute_dtype")


def _check_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda and ops.get().name == "cuda":
            raise ValueError("exoground_b200 runs on CUDA only (no CPU fallback); move inputs to the GPU")


class ViewInvariantMLP(ExgModule):
    def __init__(self, *args, **kwargs):
        super().__init__()
        self._exg_init(_take_kwargs(self, args, kwargs))
        D = self.video_embed_dim
        if D % 8:
            raise ValueError("video_embed_dim must be a multiple of 8 (16-byte rows)")
        self.tfm_modules, self.linear_layers = [], []
        self.video_pre_proj = nn.Linear(D, D, bias=False)
        self.linear_layers.append(self.video_pre_proj)
        self.ln_video_init = LayerNorm(D)
        self.mlp = nn.Sequential(nn.Linear(D, D, bias=True), nn.ReLU(), nn.Linear(D, D, bias=True))
        self.initialize_parameters()

    def initialize_parameters(self):
        """vi_encoder.py:221-232."""
        for layer in (self.video_pre_proj, self.mlp[0], self.mlp[2]):
            nn.init.normal_(layer.weight, std=0.01)
            if layer.bias is not None:
                nn.init.zeros_(layer.bias)

    def forward(self, video_embed, lang_embed=None, video_padding_mask=None, lang_padding_mask=None, audio_embed=None,
                audio_padding_mask=None, egocentric_video_embed=None, view_mask=None, interpolate_from=None):
        _check_cuda(video_embed)
        wT = self._weights()
        Tc = self.compute_dtype
        B, T, D = video_embed.shape
        x = Fn._T(video_embed.reshape(B * T, D), Tc)
        y = Fn.LinearFn.apply(x, self.video_pre_proj.weight, wT("video_pre_proj.weight"), None, Tc)
        y = Fn.LayerNormFn.apply(y, self.ln_video_init.weight, self.ln_video_init.bias, Tc)
        l0, l2 = self.mlp[0], self.mlp[2]
        z = Fn.MlpReluFn.apply(y, l0.weight, wT("mlp.0.weight"), l0.bias, l2.weight, wT("mlp.2.weight"), l2.bias,
                               torch.float32).view(B, T, D)
        return {'low_dim_features': z, 'high_dim_features': z}


class ViewInvariantEncoder(ExgModule):
    def __init__(self, *args, **kwargs):
        super().__init__()
        self._exg_init(_take_kwargs(self, args, kwargs))
        if self.feature_dim != 512:
            raise ValueError("the sm_100a kernels are specialised for feature_dim 512 / 8 heads")
        C = self.feature_dim
        self.tfm_modules = []
        self.video_unimodal_encoder = TemporalEncoder(width=C, layers=self.num_encoder_layers, heads=8)
        self.tfm_modules.append(self.video_unimodal_encoder)
        self.video_pre_proj = nn.Linear(self.video_embed_dim, C, bias=False)
        self.ln_video_init = LayerNorm(C)
        self.ln_position_init = LayerNorm(C)
        self.ln_video_post_enc = LayerNorm(C)
        if self.use_distill_nce_loss or self.use_pairwise_distill_nce_loss:
            self.exo_feature_proj = nn.Linear(C, self.video_embed_dim)
        if self.pos_enc == 'learned':
            self.temporal_pos_embed = nn.Parameter(torch.empty(1024, C))
            nn.init.normal_(self.temporal_pos_embed, std=0.01)
        elif self.pos_enc == 'sine':
            self.register_buffer('temporal_pos_embed', get_position_embedding_sine(C, 1024))
        self.initialize_parameters()

    def initialize_parameters(self):
        """vi_encoder.py:76-91."""
        nn.init.normal_(self.video_pre_proj.weight, std=0.01)
        for m in self.tfm_modules:
            proj_std = (m.width ** -0.5) * ((2 * m.layers) ** -0.5)
            attn_std, fc_std = m.width ** -0.5, (2 * m.width) ** -0.5
            for block in m.resblocks:
                nn.init.normal_(block.attn.in_proj_weight, std=attn_std)
                nn.init.normal_(block.attn.out_proj.weight, std=proj_std)
                nn.init.normal_(block.mlp.c_fc.weight, std=fc_std)
                nn.init.normal_(block.mlp.c_proj.weight, std=proj_std)

    def forward(self, video_embed, lang_embed=None, video_padding_mask=None, lang_padding_mask=None, audio_embed=None,
                audio_padding_mask=None, egocentric_video_embed=None, view_mask=None, interpolate_from=None):
        _check_cuda(video_embed)
        wT = self._weights()
        Tc = self.compute_dtype
        B, T, D = video_embed.shape
        sl = T // self.num_max_views
        x = Fn._T(video_embed.reshape(B * T, D), Tc)
        xv = Fn.LinearFn.apply(x, self.video_pre_proj.weight, wT("video_pre_proj.weight"), None, Tc)
        xv = Fn.LayerNormFn.apply(xv, self.ln_video_init.weight, self.ln_video_init.bias, torch.float32)
        p0 = int(np.random.randint(0, int(sl / 2))) if self.random_pos_start else 0      # vi_encoder.py:143-151
        rows = self.temporal_pos_embed[p0:p0 + sl]
        if self.num_max_views > 1:
            rows = rows.repeat(self.num_max_views, 1)
        pos = Fn.LayerNormFn.apply(rows.contiguous(), self.ln_position_init.weight, self.ln_position_init.bias,
                                   torch.float32)
        x0 = Fn.AddPosFn.apply(xv, pos, B, T)
        st = self.video_unimodal_encoder.run(x0, video_padding_mask.to(torch.uint8).contiguous(), B, T, wT,
                                             "video_unimodal_encoder.", self.ln_video_post_enc.weight,
                                             self.ln_video_post_enc.bias, Tc)
        v = Fn.StageMeanFn.apply(st)
        if not (self.use_distill_nce_loss and egocentric_video_embed is not None):
            raise NameError("exo_features_projected is undefined unless use_distill_nce_loss and "
                            "egocentric_video_embed are given (reference vi_encoder.py:101-104 fails the same way)")
        proj = Fn.LinearFn.apply(Fn.CastFn.apply(v, Tc), self.exo_feature_proj.weight, wT("exo_feature_proj.weight"),
                                 self.exo_feature_proj.bias, torch.float32)
        return {'low_dim_features': v.view(B, T, self.feature_dim),
                'high_dim_features': proj.view(B, T, self.video_embed_dim)}

    def compute_info_nce_loss(self, features, positive_features, temperature=0.1):
        from .exo_ground_model import info_nce_bmm
        return info_nce_bmm(features, positive_features, self.compute_dtype, 1.0 / temperature)
