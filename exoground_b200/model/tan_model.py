"""Drop-in replacement of the reference's model/tan_model.py::TemporalAligner(+Twin).

Same constructor kwargs, parameter names, forward signature and output dict (reference
tan_model.py:14-24,94-143), so `train/main.py` and `eval/eval_zeroshot_align.py` can switch by import.
Everything under forward() is libexg.so kernels driven through exoground_b200.model.engine /
functional.  Differences that are deliberate and documented in DESIGN.md:
  * the video pre-projection + LayerNorm is computed once and shared by the dual and joint branches
    (the reference recomputes it, tan_model.py:149 vs :181); the two random positional offsets are
    still drawn separately and in the reference's order;
  * `video_embed_dim` / `text_embed_dim` are constructor arguments (default 4096 as hard-coded in the
    reference, tan_model.py:42-43);
  * returned logits / features are views in stage-major storage (values and shapes identical).
"""
import numpy as np
import torch
from torch import nn
from torch.nn import LayerNorm

from .. import ops
from . import functional as Fn
from .params import ExgModule
from .tfm_model import TemporalEncoder, get_position_embedding_sine, interp_rows
from ..train.xneg import gather_rows

C = 512


class TemporalAligner(ExgModule):
    def __init__(self, num_encoder_layers=2, num_decoder_layers=2, sim='cos', language_model='word2vec',
                 pos_enc='learned', use_text_pos_enc=0, return_dual_feature=1, random_pos_start=1,
                 use_alignability_head=0, video_embed_dim=4096, text_embed_dim=4096, compute_dtype=None):
        super().__init__()
        self._exg_init(compute_dtype)
        if num_encoder_layers < 1 or num_decoder_layers < 1:
            raise ValueError("TemporalAligner needs at least one encoder and one joint layer")
        self.num_encoder_layers = num_encoder_layers
        self.num_decoder_layers = num_decoder_layers
        self.sim, self.pos_enc, self.language_model = sim, pos_enc, language_model
        self.use_text_pos_enc = use_text_pos_enc
        self.return_dual_feature = return_dual_feature
        self.xneg = None          # train.xneg.Group: all-gather the text features for cross-GPU negatives
        self.random_pos_start = random_pos_start
        self.use_alignability_head = use_alignability_head

        self.video_temporal_encoder = TemporalEncoder(width=C, layers=num_encoder_layers, heads=8)
        self.joint_temporal_encoder = TemporalEncoder(width=C, layers=num_decoder_layers, heads=8)
        self.video_pre_proj = nn.Linear(video_embed_dim, C, bias=False)
        self.text_pre_proj = nn.Linear(text_embed_dim, C, bias=False)
        self.ln_text_init = LayerNorm(C)
        self.ln_video_init = LayerNorm(C)
        self.ln_position_init = LayerNorm(C)
        self.ln_video_post_enc = LayerNorm(C)
        self.ln_joint_post_enc = LayerNorm(C)
        if pos_enc == 'learned':
            self.temporal_pos_embed = nn.Parameter(torch.empty(4096, C))
            nn.init.normal_(self.temporal_pos_embed, std=0.01)
        elif pos_enc == 'sine':
            self.register_buffer('temporal_pos_embed', get_position_embedding_sine(C, 4096))
        self.text_temporal_pos_embed = nn.Parameter(torch.empty(4096, C))
        nn.init.normal_(self.text_temporal_pos_embed, std=0.01)
        self.mlp = nn.Linear(C, C)          # unused by forward in the reference too (tan_model.py:61)
        if use_alignability_head:
            self.binary_head = nn.Linear(C, 1)
            nn.init.normal_(self.binary_head.weight, std=0.01)
            nn.init.zeros_(self.binary_head.bias)
        self.initialize_parameters()

    def initialize_parameters(self):
        """Same distributions as the reference (tan_model.py:70-91)."""
        nn.init.normal_(self.video_pre_proj.weight, std=0.01)
        nn.init.normal_(self.text_pre_proj.weight, std=0.01)
        nn.init.zeros_(self.mlp.bias)
        nn.init.normal_(self.mlp.weight, std=0.01)
        enc = self.joint_temporal_encoder
        proj_std = (enc.width ** -0.5) * ((2 * enc.layers) ** -0.5)
        attn_std, fc_std = enc.width ** -0.5, (2 * enc.width) ** -0.5
        for stack in (self.video_temporal_encoder, self.joint_temporal_encoder):
            for block in stack.resblocks:
                nn.init.normal_(block.attn.in_proj_weight, std=attn_std)
                nn.init.normal_(block.attn.out_proj.weight, std=proj_std)
                nn.init.normal_(block.mlp.c_fc.weight, std=fc_std)
                nn.init.normal_(block.mlp.c_proj.weight, std=proj_std)

    # ------------------------------------------------------------------ pieces
    def _draw(self, n):
        """Random positional offset (reference tan_model.py:157,189,218: np.random.randint(0, n/2)).  When a
        trainer has installed device-side offsets (`_pos_offsets`, a list of 0-d int64 CUDA tensors it refreshes
        before every step) they are consumed in draw order instead, so a captured CUDA graph sees new offsets."""
        offs = getattr(self, "_pos_offsets", None) if getattr(self, "_in_forward", False) else None
        if offs is not None:       # only forward() consumes them: the helper API below always draws on the host
            i = self._pos_cursor
            self._pos_cursor = i + 1
            return offs[i]
        return int(np.random.randint(0, int(n / 2))) if self.random_pos_start else 0

    def _pos_ln(self, table, p0, size, interpolate_from):
        if interpolate_from:
            rows = interp_rows(table, interpolate_from, size)
        elif torch.is_tensor(p0):      # device-side offset (CUDA-graph replay: train/step.py): rows p0 .. p0+size-1
            rows = table.index_select(0, p0 + torch.arange(size, device=table.device))
        else:
            rows = table[p0:p0 + size]
        return Fn.LayerNormFn.apply(rows.contiguous(), self.ln_position_init.weight, self.ln_position_init.bias,
                                    torch.float32)

    def _check(self, video_embed, lang_embed=None):
        if not video_embed.is_cuda and ops.get().name == "cuda":
            raise ValueError("exoground_b200 runs on CUDA only (no CPU fallback); move inputs to the GPU")
        if video_embed.dim() != 3 or (lang_embed is not None and lang_embed.dim() != 3):
            raise ValueError("expected video_embed [B,T,D] and lang_embed [B,N,D]")

    def _video_proj(self, video_embed, wT):
        """ln_video_init(video_pre_proj(v)) -> fp32 [B*T, C]  (tan_model.py:149)."""
        B, T, D = video_embed.shape
        Tc = self.compute_dtype
        x = Fn._T(video_embed.reshape(B * T, D), Tc)
        y = Fn.LinearFn.apply(x, self.video_pre_proj.weight, wT("video_pre_proj.weight"), None, Tc)
        return Fn.LayerNormFn.apply(y, self.ln_video_init.weight, self.ln_video_init.bias, torch.float32)

    def _text_proj(self, lang_embed, wT):
        """ln_text_init(text_pre_proj(t)) -> fp32 [B*N, C]  (tan_model.py:225-228)."""
        B, N, D = lang_embed.shape
        Tc = self.compute_dtype
        x = Fn._T(lang_embed.reshape(B * N, D), Tc)
        y = Fn.LinearFn.apply(x, self.text_pre_proj.weight, wT("text_pre_proj.weight"), None, Tc)
        return Fn.LayerNormFn.apply(y, self.ln_text_init.weight, self.ln_text_init.bias, torch.float32)

    def _video_stages(self, xv, B, T, vmask_u8, p0, interpolate_from, wT):
        pos = self._pos_ln(self.temporal_pos_embed, p0, T, interpolate_from)
        x0 = Fn.AddPosFn.apply(xv, pos, B, T)
        return self.video_temporal_encoder.run(x0, vmask_u8, B, T, wT, "video_temporal_encoder.",
                                               self.ln_video_post_enc.weight, self.ln_video_post_enc.bias,
                                               self.compute_dtype)            # [E, B*T, C]

    def _joint_stages(self, xv, xt, B, T, N, vmask_u8, tmask_u8, p0, interpolate_from, wT):
        pos = self._pos_ln(self.temporal_pos_embed, p0, T, interpolate_from)
        x0 = Fn.JointConcatFn.apply(xv, pos, xt, B, T, N)
        mask = torch.cat((vmask_u8, tmask_u8), dim=1).contiguous() if vmask_u8 is not None else None
        return self.joint_temporal_encoder.run(x0, mask, B, T + N, wT, "joint_temporal_encoder.",
                                               self.ln_joint_post_enc.weight, self.ln_joint_post_enc.bias,
                                               self.compute_dtype)            # [D, B*(T+N), C]

    def _text_with_time(self, xt, B, N, p0, interpolate_from):
        pos = self._pos_ln(self.text_temporal_pos_embed, p0, N, interpolate_from)
        return Fn.AddPosFn.apply(xt, pos, B, N)

    def _head(self, x_rows, wT):
        return Fn.RowDotFn.apply(x_rows, self.binary_head.weight, self.binary_head.bias)

    # ------------------------------------------------------------------ forward
    def forward(self, video_embed, lang_embed, video_padding_mask, lang_padding_mask, text_timestamp=None,
                interpolate_from=None, abs_text_pos=None):
        self._check(video_embed, lang_embed)
        self._pos_cursor = 0
        self._in_forward = True
        try:
            return self._forward(video_embed, lang_embed, video_padding_mask, lang_padding_mask, interpolate_from)
        finally:
            self._in_forward = False

    def _forward(self, video_embed, lang_embed, video_padding_mask, lang_padding_mask, interpolate_from):
        wT = self._weights()
        Tc = self.compute_dtype
        B, T, _ = video_embed.shape
        N = lang_embed.shape[1]
        E, D = self.num_encoder_layers, self.num_decoder_layers
        vmask = video_padding_mask.to(torch.uint8).contiguous()
        tmask = lang_padding_mask.to(torch.uint8).contiguous()
        # random positional offsets in the reference's draw order (tan_model.py:157,218,189)
        pv = 0 if interpolate_from else self._draw(T)
        pt = 0 if (interpolate_from or not self.use_text_pos_enc) else self._draw(N)
        pj = 0 if interpolate_from else self._draw(T)

        xv = self._video_proj(video_embed, wT)                                     # fp32 [B*T, C]
        xt = self._text_proj(lang_embed, wT)                                       # fp32 [B*N, C]
        xg = self.xneg if self.xneg is not None and self.xneg.world > 1 else None
        Bg = B * (xg.world if xg is not None else 1)
        xt_time = self._text_with_time(xt, B, N, pt, interpolate_from) if self.use_text_pos_enc else xt

        # The joint branch only needs the shared projections (tan_model.py:176-203), not the video encoder's output,
        # so a trainer may give it a stream of its own (`_branch_stream`): its latency-bound kernels (attention at
        # S <= 128, LayerNorm, N = 512 GEMMs) then fill the SMs the other branch leaves idle, in forward and - since
        # autograd replays each node on its forward stream - in backward.
        side = getattr(self, "_branch_stream", None) if xv.is_cuda and xg is None else None
        main = torch.cuda.current_stream() if side is not None else None

        def joint_branch():
            st_j = self._joint_stages(xv, xt_time, B, T, N, vmask, tmask, pj, interpolate_from, wT)
            jvn, jtn = Fn.SplitL2NormFn.apply(st_j, B, T, N)                       # [D,B*T,C], [D,B*N,C]
            lj = Fn.SimLogitsFn.apply(jvn, gather_rows(jtn, xg))                   # [D, B*T, Bg*N]
            ja = None
            if self.use_alignability_head:
                jt = Fn.GatherRowsFn.apply(st_j, D * B * N, (N, T + N, T), Tc)       # joint_text_out rows
                ja = self._head(jt, wT).view(D, B, N, 1).permute(1, 0, 2, 3)
            return lj, ja

        if side is not None:
            side.wait_stream(main)
            with torch.cuda.stream(side):
                logits_joint, joint_align = joint_branch()
        st_v = self._video_stages(xv, B, T, vmask, pv, interpolate_from, wT)       # [E, B*T, C]
        vn = Fn.L2NormFn.apply(st_v, Tc)                                           # [E, B*T, C]
        tn = Fn.L2NormFn.apply(xt, Tc)                                             # [B*N, C]
        # cross-GPU negatives (train/xneg.py): columns = the text features of ALL ranks, rows stay local
        logits_dual = Fn.SimLogitsFn.apply(vn, gather_rows(tn, xg))                # [E, B*T, Bg*N]
        if side is not None:
            main.wait_stream(side)
            if not torch.cuda.is_current_stream_capturing():
                for t in (logits_joint, joint_align):
                    if t is not None:
                        t.record_stream(main)
        else:
            logits_joint, joint_align = joint_branch()

        out = {'logits_dual': logits_dual.view(E, B, T, Bg, N).permute(1, 0, 2, 3, 4),
               'logits_joint': logits_joint.view(D, B, T, Bg, N).permute(1, 0, 2, 3, 4)}
        if self.return_dual_feature:
            out['dual_feature_video'] = vn.view(E, B, T, C).permute(1, 0, 2, 3)
            out['dual_feature_text'] = tn.view(B, N, C)
        if self.use_alignability_head:
            out['dual_logits_alignability'] = self._head(xt, wT).view(B, N, 1)
            out['joint_logits_alignability'] = joint_align
        return out

    # ------------------------------------------------------------------ reference helper API
    def get_visual_feature(self, video_embed, video_padding_mask, interpolate_from=None):
        """[B,E,T,C] stage features of the video encoder (tan_model.py:146-173)."""
        self._check(video_embed)
        wT = self._weights()
        B, T, _ = video_embed.shape
        p0 = 0 if interpolate_from else self._draw(T)
        st = self._video_stages(self._video_proj(video_embed, wT), B, T,
                                video_padding_mask.to(torch.uint8).contiguous(), p0, interpolate_from, wT)
        return st.view(self.num_encoder_layers, B, T, C).permute(1, 0, 2, 3)

    def get_textual_feature(self, lang_embed):
        """[B,N,C] (tan_model.py:225-228)."""
        self._check(lang_embed)
        B, N, _ = lang_embed.shape
        return self._text_proj(lang_embed, self._weights()).view(B, N, C)

    def get_textual_feature_with_time(self, lang_embed, text_timestamp=None, interpolate_from=None):
        self._check(lang_embed)
        B, N, _ = lang_embed.shape
        xt = self._text_proj(lang_embed, self._weights())
        p0 = 0 if interpolate_from else self._draw(N)
        return self._text_with_time(xt, B, N, p0, interpolate_from).view(B, N, C)

    def get_joint_feature(self, video_embed, video_padding_mask, lang_embed_with_time, lang_padding_mask,
                          interpolate_from=None):
        """([B,D,T,C], [B,D,N,C]) joint-encoder stage features (tan_model.py:176-203).
        lang_embed_with_time is the already projected text [B,N,512]."""
        self._check(video_embed)
        wT = self._weights()
        B, T, _ = video_embed.shape
        N = lang_embed_with_time.shape[1]
        D = self.num_decoder_layers
        p0 = 0 if interpolate_from else self._draw(T)
        xt = lang_embed_with_time.reshape(B * N, C).float().contiguous()
        st = self._joint_stages(self._video_proj(video_embed, wT), xt, B, T, N,
                                video_padding_mask.to(torch.uint8).contiguous(),
                                lang_padding_mask.to(torch.uint8).contiguous(), p0, interpolate_from, wT)
        st = st.view(D, B, T + N, C).permute(1, 0, 2, 3)
        return st[:, :, :T], st[:, :, T:]

    @staticmethod
    def _split_interp(interpolate_from):
        if isinstance(interpolate_from, (list, tuple)):
            assert len(interpolate_from) == 2
            return interpolate_from[0], interpolate_from[1]
        return interpolate_from, None

    def _joint_sim_parts(self, video_embed, lang_embed, interpolate_from):
        interpolate_from, text_from = self._split_interp(interpolate_from)
        wT = self._weights()
        B, T, _ = video_embed.shape
        N = lang_embed.shape[1]
        xt = self._text_proj(lang_embed, wT)
        if self.use_text_pos_enc:
            xt = self._text_with_time(xt, B, N, 0 if text_from else self._draw(N), text_from)
        p0 = 0 if interpolate_from else self._draw(T)
        st = self._joint_stages(self._video_proj(video_embed, wT), xt, B, T, N, None, None, p0, interpolate_from, wT)
        return st, B, T, N

    def get_text_visual_sim_joint(self, video_embed, lang_embed, interpolate_from=None):
        """[B,S,T,K] per-sample joint logits, zero padding masks (tan_model.py:231-258)."""
        self._check(video_embed, lang_embed)
        st, B, T, N = self._joint_sim_parts(video_embed, lang_embed, interpolate_from)
        D = self.num_decoder_layers
        jvn, jtn = Fn.SplitL2NormFn.apply(st, B, T, N)
        # per-sample products: batch = D*B, [T,C] x [N,C]^T
        lg = Fn.SimLogitsFn.apply(jvn.view(D * B, T, C), jtn.view(D * B, N, C))
        return lg.view(D, B, T, N).permute(1, 0, 2, 3)

    def get_text_visual_sim_dual(self, video_embed, lang_embed, interpolate_from=None):
        """[B,S,T,K] per-sample dual logits (tan_model.py:261-277)."""
        self._check(video_embed, lang_embed)
        wT = self._weights()
        Tc = self.compute_dtype
        B, T, _ = video_embed.shape
        N = lang_embed.shape[1]
        E = self.num_encoder_layers
        tn = Fn.L2NormFn.apply(self._text_proj(lang_embed, wT), Tc).view(B, N, C)
        p0 = 0 if interpolate_from else self._draw(T)
        st = self._video_stages(self._video_proj(video_embed, wT), B, T, None, p0, interpolate_from, wT)
        vn = Fn.L2NormFn.apply(st, Tc).view(E, B, T, C)
        tn_rep = tn.unsqueeze(0).expand(E, B, N, C).reshape(E * B, N, C)                 # small copy
        lg = Fn.SimLogitsFn.apply(vn.view(E * B, T, C), tn_rep)                         # [E*B,T,N]
        return lg.view(E, B, T, N).permute(1, 0, 2, 3)

    def get_alignability(self, video_embed, lang_embed, interpolate_from=None):
        """{'alignability-dual' [B,N,1], 'alignability-joint' [B,D,N,1]} (tan_model.py:280-306)."""
        self._check(video_embed, lang_embed)
        st, B, T, N = self._joint_sim_parts(video_embed, lang_embed, interpolate_from)
        wT = self._weights()
        Tc = self.compute_dtype
        D = self.num_decoder_layers
        jt = Fn.GatherRowsFn.apply(st, D * B * N, (N, T + N, T), Tc)
        dual = self._head(self._text_proj(lang_embed, wT), wT).view(B, N, 1)
        return {'alignability-dual': dual,
                'alignability-joint': self._head(jt, wT).view(D, B, N, 1).permute(1, 0, 2, 3)}


class TwinTemporalAligner(nn.Module):
    """Online / EMA-target pair (reference tan_model.py:309-345; the reference class cannot be constructed
    as shipped — it references a removed `bert` attribute — so this is the working equivalent).
    `_momentum_update` is one fused kernel over the flat parameter buffers."""

    def __init__(self, m=0.999, *args, **kwargs):
        super().__init__()
        self.m = m
        self.online = TemporalAligner(*args, **kwargs)
        self.target = TemporalAligner(*args, **kwargs)
        self._copy_param()
        for name in ("get_visual_feature", "get_joint_feature", "get_textual_feature_with_time",
                     "get_textual_feature", "get_text_visual_sim_joint", "get_text_visual_sim_dual",
                     "get_alignability"):
            setattr(self, name, getattr(self.online, name))
        self.target.random_pos_start = 0

    def _copy_param(self):
        for po, pt in zip(self.online.parameters(), self.target.parameters()):
            pt.data.copy_(po.data)
            pt.requires_grad = False

    @torch.no_grad()
    def _momentum_update(self):
        fo = self.online.flat_params()
        # a trainer moves the online model's never-used parameters to the tail of its flat buffer
        # (FlatParams.set_unused): give the target the same layout before updating buffer against buffer
        self.target._flat.set_unused(fo.unused)
        ft = self.target.flat_params()
        if ft.index != fo.index:
            raise RuntimeError("TwinTemporalAligner: online / target flat layouts differ")
        ops.get().ema(ft.flat, fo.flat, self.m)
        self.target._shadow_stale = True

    def forward(self, *args, **kwargs):
        return self.online(*args, **kwargs)

    @torch.no_grad()
    def forward_from_ema(self, *args, **kwargs):
        return self.target(*args, **kwargs)
