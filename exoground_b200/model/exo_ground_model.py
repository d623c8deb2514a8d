"""Drop-in replacement of the reference's model/exo_ground_model.py::ExoGroundingTransformer (the `joint` /
`grounding` model of train/main_egoexo4d_distributed.py:626-713): video and text unimodal encoders -> stage
mean -> multimodal encoder -> cross-attention decoder -> Linear(512, 2) interval regression, plus the optional
512 -> D_v projection used by the ego-distillation InfoNCE.  Same constructor kwargs, parameter names, forward
signature and output dict (reference exo_ground_model.py:13-31,127-183).  Everything under forward() is
libexg.so kernels driven through exoground_b200.model.engine / functional.

Audio branch (`use_audio`, reference :135-139,347-350): `ln_audio_init(audio_pre_proj(audio))` + position rows is
inserted between the video and the text rows of the joint sequence.  As in the reference it only works with
`use_decoder=False`: with a decoder the reference hands 2T memory rows and a T-wide key-padding mask to
nn.MultiheadAttention, which asserts (:171); the same AssertionError is raised here.

Pairwise multi-view InfoNCE (:236-292): the reference selects the valid (clip, second) rows of two view chunks,
gives each a sequence axis of length ONE and takes log_softmax over that axis, so every pair contributes exactly
0 and exactly zero gradient (verified by running the reference: oracle/make_golden.py::pairwise_probe).  The
drop-in reproduces that value and gradient (`pairwise_info_nce`), including the 4-D bmm failure when a chunk
holds a single second.
"""
import numpy as np
import torch
from torch import nn
from torch.nn import LayerNorm

from .. import ops
from . import functional as Fn
from .params import ExgModule
from .tfm_model import TemporalEncoder, TemporalDecoder, get_position_embedding_sine, interp_rows

INV_TEMP_DISTILL = 1.0 / 0.1     # compute_info_nce_loss(..., temperature=0.1)


class ExoGroundingTransformer(ExgModule):
    def __init__(self, num_encoder_layers=2, num_decoder_layers=2, use_decoder=True, sim='cos', pos_enc='learned',
                 use_text_pos_enc=0, random_pos_start=1, use_audio=False, video_embed_dim=4096, text_embed_dim=4096,
                 audio_embed_dim=2304, feature_dim=512, use_distill_nce_loss=False, multi_view=False,
                 num_max_views=1, use_pairwise_distill_nce_loss=False, pairwise_distill_mode="all",
                 compute_dtype=None):
        super().__init__()
        self._exg_init(compute_dtype)
        if feature_dim != 512:
            raise ValueError("the sm_100a kernels are specialised for feature_dim 512 / 8 heads")
        self.num_encoder_layers, self.num_decoder_layers = num_encoder_layers, num_decoder_layers
        self.use_decoder, self.sim, self.pos_enc = use_decoder, sim, pos_enc
        self.use_text_pos_enc, self.random_pos_start, self.use_audio = use_text_pos_enc, random_pos_start, use_audio
        self.text_embed_dim, self.audio_embed_dim, self.video_embed_dim = text_embed_dim, audio_embed_dim, video_embed_dim
        self.feature_dim = feature_dim
        self.use_distill_nce_loss, self.multi_view, self.num_max_views = use_distill_nce_loss, multi_view, num_max_views
        self.use_pairwise_distill_nce_loss, self.pairwise_distill_mode = use_pairwise_distill_nce_loss, pairwise_distill_mode

        self.tfm_modules = []
        self.multi_modal_encoder = TemporalEncoder(width=feature_dim, layers=num_encoder_layers, heads=8)
        self.tfm_modules.append(self.multi_modal_encoder)
        if use_decoder:
            self.decoder = TemporalDecoder(width=feature_dim, layers=num_decoder_layers, heads=8)
            self.tfm_modules.append(self.decoder)
        self.grounding_head = nn.Linear(feature_dim, 2)
        self.video_unimodal_encoder = TemporalEncoder(width=feature_dim, layers=num_encoder_layers, heads=8)
        self.tfm_modules.append(self.video_unimodal_encoder)
        self.text_unimodal_encoder = TemporalEncoder(width=feature_dim, layers=num_encoder_layers, heads=8)
        self.tfm_modules.append(self.text_unimodal_encoder)
        self.video_pre_proj = nn.Linear(video_embed_dim, feature_dim, bias=False)
        self.text_pre_proj = nn.Linear(text_embed_dim, feature_dim, bias=False)
        self.ln_text_init = LayerNorm(feature_dim)
        self.ln_video_init = LayerNorm(feature_dim)
        self.ln_position_init = LayerNorm(feature_dim)
        self.ln_joint_post_enc = LayerNorm(feature_dim)
        self.ln_video_post_enc = LayerNorm(feature_dim)
        self.ln_text_post_enc = LayerNorm(feature_dim)
        if use_distill_nce_loss or use_pairwise_distill_nce_loss:
            self.exo_feature_proj = nn.Linear(feature_dim, video_embed_dim)
        if use_audio:
            self.ln_audio_init = LayerNorm(feature_dim)
            self.audio_pre_proj = nn.Linear(audio_embed_dim, feature_dim, bias=False)
        if pos_enc == 'learned':
            self.temporal_pos_embed = nn.Parameter(torch.empty(1024, feature_dim))
            nn.init.normal_(self.temporal_pos_embed, std=0.01)
        elif pos_enc == 'sine':
            self.register_buffer('temporal_pos_embed', get_position_embedding_sine(feature_dim, 1024))
        self.text_temporal_pos_embed = nn.Parameter(torch.empty(text_embed_dim, feature_dim))   # sic (:99)
        nn.init.normal_(self.text_temporal_pos_embed, std=0.01)
        self.mlp = nn.Linear(feature_dim, feature_dim)      # unused by forward in the reference too (:102)
        self.initialize_parameters()

    def initialize_parameters(self):
        """Same distributions as the reference (exo_ground_model.py:106-124)."""
        layers = [self.video_pre_proj, self.text_pre_proj, self.mlp, self.grounding_head]
        if self.use_audio:
            layers.append(self.audio_pre_proj)
        for layer in layers:
            nn.init.normal_(layer.weight, std=0.01)
            if layer.bias is not None:
                nn.init.zeros_(layer.bias)
        for m in self.tfm_modules:
            proj_std = (m.width ** -0.5) * ((2 * m.layers) ** -0.5)
            attn_std, fc_std = m.width ** -0.5, (2 * m.width) ** -0.5
            for block in m.resblocks:
                nn.init.normal_(block.attn.in_proj_weight, std=attn_std)
                nn.init.normal_(block.attn.out_proj.weight, std=proj_std)
                nn.init.normal_(block.mlp.c_fc.weight, std=fc_std)
                nn.init.normal_(block.mlp.c_proj.weight, std=proj_std)

    # ------------------------------------------------------------------ pieces
    def _check(self, *ts):
        for t in ts:
            if t is not None and not t.is_cuda and ops.get().name == "cuda":
                raise ValueError("exoground_b200 runs on CUDA only (no CPU fallback); move inputs to the GPU")

    def _draw(self, n):
        """Random positional offset (exo_ground_model.py:305,338,377: np.random.randint(0, n/2)).  A trainer that
        replays a captured CUDA graph installs device-side offsets (`_pos_offsets`: 0-d int64 CUDA tensors it refreshes
        before every replay), consumed in draw order inside forward() only."""
        offs = getattr(self, "_pos_offsets", None) if getattr(self, "_in_forward", False) else None
        if offs is not None and self.random_pos_start:
            i = self._pos_cursor
            self._pos_cursor = i + 1
            return offs[i]
        return int(np.random.randint(0, int(n / 2))) if self.random_pos_start else 0

    def _pos_ln(self, table, p0, size, interpolate_from, repeat=1):
        if interpolate_from:
            rows = interp_rows(table, interpolate_from, size)
        elif torch.is_tensor(p0):      # device-side offset (graph replay): rows p0 .. p0 + size - 1
            rows = table.index_select(0, p0 + torch.arange(size, device=table.device))
        else:
            rows = table[p0:p0 + size]
        if repeat > 1:
            rows = rows.repeat(repeat, 1)
        return Fn.LayerNormFn.apply(rows.contiguous(), self.ln_position_init.weight, self.ln_position_init.bias,
                                    torch.float32)

    def _proj_ln(self, x, lin, ln, wT, name):
        B, S, D = x.shape
        xt = Fn._T(x.reshape(B * S, D), self.compute_dtype)
        y = Fn.LinearFn.apply(xt, lin.weight, wT(name + ".weight"), None, self.compute_dtype)
        return Fn.LayerNormFn.apply(y, ln.weight, ln.bias, torch.float32)          # fp32 [B*S, C]

    def _text_proj(self, lang_embed, wT):
        return self._proj_ln(lang_embed, self.text_pre_proj, self.ln_text_init, wT, "text_pre_proj")

    def _video_unimodal_mean(self, video_embed, vmask_u8, wT, interpolate_from=None):
        """get_unimodal_features('video', ...).mean(dim=1) (exo_ground_model.py:294-321,153) -> fp32 [B*T, C]."""
        B, T, _ = video_embed.shape
        sl = T // self.num_max_views
        xv = self._proj_ln(video_embed, self.video_pre_proj, self.ln_video_init, wT, "video_pre_proj")
        p0 = 0 if interpolate_from else self._draw(sl)
        pos = self._pos_ln(self.temporal_pos_embed, p0, sl, interpolate_from, self.num_max_views)
        x0 = Fn.AddPosFn.apply(xv, pos, B, T)
        st = self.video_unimodal_encoder.run(x0, vmask_u8, B, T, wT, "video_unimodal_encoder.",
                                             self.ln_video_post_enc.weight, self.ln_video_post_enc.bias,
                                             self.compute_dtype)
        return Fn.StageMeanFn.apply(st)

    def _text_unimodal_mean(self, xt, tmask_u8, B, N, wT):
        st = self.text_unimodal_encoder.run(xt, tmask_u8, B, N, wT, "text_unimodal_encoder.",
                                            self.ln_text_post_enc.weight, self.ln_text_post_enc.bias,
                                            self.compute_dtype)
        return Fn.StageMeanFn.apply(st)

    def _audio_proj(self, audio_embed, wT):
        """get_audio_feature (reference :390-393) -> fp32 [B*T, C]."""
        if not self.use_audio:
            raise AttributeError("'ExoGroundingTransformer' object has no attribute 'audio_pre_proj'")
        return self._proj_ln(audio_embed, self.audio_pre_proj, self.ln_audio_init, wT, "audio_pre_proj")

    def _unimodal_pair(self, video_embed, vmask_u8, xt, tmask_u8, B, N, wT):
        """The two unimodal encoders are independent (exo_ground_model.py:153-154): with a trainer-provided
        `_branch_stream` the (short, latency-bound) text encoder runs beside the video encoder, forward and - through
        autograd's stream replay - backward.  np.random draw order is the reference's (video first)."""
        side = getattr(self, "_branch_stream", None) if xt.is_cuda else None
        if side is None:
            v_enc = self._video_unimodal_mean(video_embed, vmask_u8, wT)     # reference passes no interpolate_from here
            return v_enc, self._text_unimodal_mean(xt, tmask_u8, B, N, wT)
        main = torch.cuda.current_stream()
        side.wait_stream(main)
        with torch.cuda.stream(side):
            t_enc = self._text_unimodal_mean(xt, tmask_u8, B, N, wT)
        v_enc = self._video_unimodal_mean(video_embed, vmask_u8, wT)
        main.wait_stream(side)
        if not torch.cuda.is_current_stream_capturing():
            t_enc.record_stream(main)
        return v_enc, t_enc

    def _grounding(self, v_enc, t_enc, vmask_u8, tmask_u8, B, T, N, wT, a_feat=None, amask_u8=None,
                   interpolate_from=None):
        """multi-modal encoder -> decoder -> grounding head (exo_ground_model.py:161-177, 323-362).  With audio the
        joint sequence is [video | audio + ln_position_init(pos) | text] (:347-350)."""
        Tc = self.compute_dtype
        E = self.num_encoder_layers
        sl = T // self.num_max_views
        p0 = 0 if interpolate_from else self._draw(sl)                 # get_joint_feature draws even without audio (:338)
        if a_feat is not None:
            assert a_feat.shape == v_enc.shape, \
                "Audio and video inputs must match in all dimensions (batch size and timesteps and feature size)"
            if self.use_decoder:       # reference :171 -> F.multi_head_attention_forward's key_padding_mask check
                raise AssertionError(f"Expected key_padded_mask.shape[1] to be {2 * T}, but got {T}")
            pos = self._pos_ln(self.temporal_pos_embed, p0, sl, interpolate_from, self.num_max_views)
            x0 = Fn.JointConcat3Fn.apply(v_enc, a_feat, pos, t_enc, B, T, N)
            mask = torch.cat((vmask_u8, amask_u8, tmask_u8), dim=1).contiguous()
            Tctx = 2 * T
        else:
            x0 = Fn.JointConcatFn.apply(v_enc, None, t_enc, B, T, N)
            mask = torch.cat((vmask_u8, tmask_u8), dim=1).contiguous()
            Tctx = T
        S = Tctx + N
        st = self.multi_modal_encoder.run(x0, mask, B, S, wT, "multi_modal_encoder.",
                                          self.ln_joint_post_enc.weight, self.ln_joint_post_enc.bias, Tc)
        head_w, head_b = self.grounding_head.weight, self.grounding_head.bias
        if self.use_decoder:
            last = st[E - 1]                                                       # [B*(T+N), C]
            txt = Fn.GatherRowsFn.apply(last, B * N, (N, S, Tctx), torch.float32)   # decoder input x (fp32 stream)
            ctx = Fn.GatherRowsFn.apply(last, B * T, (T, S, 0), Tc)                 # decoder memory
            dec = self.decoder.run(txt, ctx, tmask_u8, vmask_u8, B, N, T, wT, "decoder.", Tc)
            g = Fn.LinearFn.apply(dec, head_w, wT("grounding_head.weight"), head_b, torch.float32)
            return g.view(B, N, 2)
        # no decoder: the head is applied to the text rows of ALL stages (reference :177) -> [B,E,N,2]
        txt = Fn.GatherRowsFn.apply(st, E * B * N, (N, S, Tctx), Tc)
        g = Fn.LinearFn.apply(txt, head_w, wT("grounding_head.weight"), head_b, torch.float32)
        return g.view(E, B, N, 2).permute(1, 0, 2, 3)

    def _exo_proj(self, v_enc, B, T, wT):
        proj = Fn.LinearFn.apply(Fn.CastFn.apply(v_enc, self.compute_dtype), self.exo_feature_proj.weight,
                                 wT("exo_feature_proj.weight"), self.exo_feature_proj.bias, torch.float32)
        return proj.view(B, T, self.video_embed_dim)

    # ------------------------------------------------------------------ forward
    def forward(self, video_embed, lang_embed, video_padding_mask, lang_padding_mask, audio_embed=None,
                audio_padding_mask=None, egocentric_video_embed=None, view_mask=None, interpolate_from=None):
        self._check(video_embed, lang_embed, audio_embed)
        self._pos_cursor = 0
        self._in_forward = True
        try:
            return self._forward(video_embed, lang_embed, video_padding_mask, lang_padding_mask, audio_embed,
                                 audio_padding_mask, egocentric_video_embed, view_mask, interpolate_from)
        finally:
            self._in_forward = False

    def _forward(self, video_embed, lang_embed, video_padding_mask, lang_padding_mask, audio_embed,
                 audio_padding_mask, egocentric_video_embed, view_mask, interpolate_from):
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
        v_enc, t_enc = self._unimodal_pair(video_embed, vmask, xt, tmask, B, N, wT)
        proj = None
        if (self.use_distill_nce_loss and egocentric_video_embed is not None) or \
                (self.multi_view and self.use_pairwise_distill_nce_loss):
            proj = self._exo_proj(v_enc, B, T, wT)
        grounding = self._grounding(v_enc, t_enc, vmask, tmask, B, T, N, wT, a_feat, amask, interpolate_from)
        return self._outputs(grounding, v_enc.view(B, T, self.feature_dim), proj, egocentric_video_embed)

    def _outputs(self, grounding, v_enc, proj, ego):
        out = {'interval_preds': grounding, 'low_dim_features': v_enc}
        if self.use_distill_nce_loss or self.use_pairwise_distill_nce_loss:
            if proj is None:
                raise NameError("exo_features_projected is undefined without egocentric_video_embed "
                                "(reference exo_ground_model.py:156-159,180-181 raises the same way)")
            out['high_dim_features'] = proj
        return out

    def get_low_dim_target_features(self, video_embed, video_padding_mask):
        """reference :185-188."""
        self._check(video_embed)
        B, T, _ = video_embed.shape
        v = self._video_unimodal_mean(video_embed, video_padding_mask.to(torch.uint8).contiguous(), self._weights())
        return {'low_dim_features': v.view(B, T, self.feature_dim)}

    def compute_info_nce_loss(self, features, positive_features, temperature=0.1):
        """reference :206-234: per-clip T x T InfoNCE between two [B,T,D] sequences."""
        return info_nce_bmm(features, positive_features, self.compute_dtype, 1.0 / temperature)

    def compute_pairwise_info_nce_loss(self, features, temperature=0.1, view_mask=None):
        """reference :236-292."""
        return pairwise_info_nce(features, view_mask, self.num_max_views)


class _ZeroOfFn(torch.autograd.Function):
    """A 0-d zero that stays attached to the graph of `x` with an exactly-zero gradient."""

    @staticmethod
    def forward(ctx, x):
        ctx.meta = (x.shape, x.dtype)
        return torch.zeros((), dtype=torch.float32, device=x.device)

    @staticmethod
    def backward(ctx, g):
        shape, dtype = ctx.meta
        return torch.zeros(shape, dtype=dtype, device=g.device)


def pairwise_info_nce(features, view_mask, num_max_views):
    """compute_pairwise_info_nce_loss (exo_ground_model.py:236-292, keystep_ground_model.py:251-307).

    features [B, V*T', D] and view_mask [B, V*T'] are chunked into V views; for each view pair the reference
    gathers the rows valid in both views ([M, D]), unsqueezes a length-1 sequence axis and evaluates the
    InfoNCE over that axis: bmm -> [M,1,1], log_softmax over one element = 0.  The loss of every pair is therefore
    0 and so is its gradient; what remains observable is (a) the result is attached to the graph iff some pair has
    a valid row, (b) the 4-D bmm RuntimeError when T' == 1, (c) the chunking errors for a missing mask.  One
    device->host read (`any`) like the reference's `valid_mask.any()`."""
    if view_mask is None:
        raise TypeError("chunk(): argument 'input' must be Tensor, not NoneType")
    V = num_max_views
    fchunks = torch.chunk(features, V, dim=1)
    mchunks = torch.chunk(view_mask, V, dim=1)
    if len(mchunks) < V or len(fchunks) < V:
        raise IndexError("tuple index out of range")
    if fchunks[0].shape[1] == 1:
        raise RuntimeError("batch1 must be a 3D tensor")
    m = torch.stack([c.bool() for c in mchunks[:V]])                    # [V, B, T']
    both = (m.unsqueeze(0) & m.unsqueeze(1)).flatten(2).any(-1)         # [V, V]
    iu = torch.triu_indices(V, V, offset=1, device=both.device)
    if V > 1 and bool(both[iu[0], iu[1]].any()):
        return _ZeroOfFn.apply(features)
    return torch.tensor(0.0).to(features.device)


def info_nce_bmm(f1, f2, compute_dtype, inv_temp=INV_TEMP_DISTILL):
    """-mean log_softmax(normalize(f1) normalize(f2)^T / temp)[diag] over [B,T,D] inputs (F.normalize's eps only
    matters for all-zero rows)."""
    assert f1.size(1) == f2.size(1)
    a = Fn.L2NormFn.apply(f1.contiguous(), compute_dtype)
    b = Fn.L2NormFn.apply(f2.contiguous(), compute_dtype)
    return Fn.InfoNceBmmFn.apply(a, b, inv_temp)
