"""One ExoGround / keystep-grounding training step on one GPU (one process per GPU): the body of the reference's
train() loop (train/main_egoexo4d_distributed.py:105-222: forward :145-153, get_loss :165-168, backward :175,
clip + optimizer step :176-181) with DistributedDataParallel (:611-614) replaced by ONE NCCL all-reduce of the flat
gradient buffer and torch.optim.AdamW by the fused kernel (train/flat.py).  The grounding / view-invariant losses
have no cross-sample terms, so the data-parallel mode is DDP-equivalent: local loss, mean gradient.

The step runs eagerly: the reference loss dictionary carries per-camera-rank host statistics (one read-back per
loss here instead of one `.item()` per element), which keeps it out of a CUDA graph."""
import torch

from .. import ops
from . import loss_egoexo4d as exo_loss
from .flat import FlatAdamW

_DEVICE_KEYS = ("video_features", "narration_features", "video_padding_mask", "narration_padding_mask",
                "ego_video_features", "view_rank_label", "view_rank_neg_label", "same_view_neg_idxs",
                "starts", "ends", "mean", "duration", "view_available_mask")


class ExoTrainStep:
    def __init__(self, model, args, lr=1e-4, wd=1e-5, betas=(0.9, 0.999), eps=1e-8, clip_grad=0.0, device="cuda",
                 process_group=None, branch_streams=True):
        if ops.get().name == "cuda" and not (torch.cuda.is_available() and torch.device(device).type == "cuda"):
            raise RuntimeError("exoground_b200 training needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        self.model, self.args, self.clip_grad = model.to(device), args, clip_grad
        self.opt = FlatAdamW(self.model, lr, wd, betas, eps, process_group)
        self.h2d_bytes = 0
        self._copy_stream, self._staged = None, None
        # text unimodal encoder beside the video unimodal encoder (model/exo_ground_model.py::_unimodal_pair)
        self.branch_stream = None
        if branch_streams and self.device.type == "cuda" and ops.get().name == "cuda":
            self.branch_stream = torch.cuda.Stream(device=self.device)
            self.model._branch_stream = self.branch_stream
            quiet = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
            if quiet is not None:
                quiet(False)

    def to_device(self, batch):
        out, n = dict(batch), 0
        for k in _DEVICE_KEYS:
            v = batch.get(k)
            if torch.is_tensor(v):
                out[k] = v.to(self.device, non_blocking=True)
                n += v.numel() * v.element_size()
        self.h2d_bytes = n
        return out

    def prefetch(self, batch):
        """Starts the H2D copies of a FUTURE batch on a copy stream (pinned host tensors make them asynchronous), so
        they overlap the step that is running; step(batch) of the same dict object picks the staged copy up.  The
        collate dict carries [B,V,T,D] ego features (251 MB at B=16, T=1024): un-overlapped that copy is a third of
        the step."""
        if self.device.type != "cuda":
            return
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        cs = self._copy_stream
        with torch.cuda.stream(cs):
            staged = self.to_device(batch)
            ev = torch.cuda.Event()
            ev.record(cs)
        self._staged = (id(batch), staged, ev)

    def _take(self, batch):
        if self._staged is not None and self._staged[0] == id(batch):
            _, staged, ev = self._staged
            self._staged = None
            main = torch.cuda.current_stream()
            main.wait_event(ev)
            for v in staged.values():
                if torch.is_tensor(v) and v.is_cuda:
                    v.record_stream(main)
            return staged
        return self.to_device(batch)

    def _fwd_bwd(self, b):
        m, args = self.model, self.args
        ego = b.get("ego_video_features") if getattr(args, "use_distill_nce_loss", False) else None
        kw = {}
        if ego is not None:
            kw["egocentric_video_embed"] = ego[:, 0] if ego.dim() == 4 else ego
        out = m(b["video_features"], b["narration_features"], b["video_padding_mask"].bool(),
                b["narration_padding_mask"].bool(), **kw)
        L, iou = exo_loss.get_loss(b, out, b["narration_padding_mask"], args)
        L["loss"].backward()
        if self.branch_stream is not None:
            torch.cuda.current_stream().wait_stream(self.branch_stream)
        return L, iou

    def step(self, batch, next_batch=None):
        """Host batch (the collate dict of data/loader_egoexo4d.py:567-598) in, loss dict out.  `next_batch`
        (optional) is copied to the device on the copy stream while this step computes."""
        self.model.train()
        b = self._take(batch)
        if not self.opt.attached:
            for p in self.model.parameters():
                p.grad = None
            self._fwd_bwd(b)                    # discovery pass: which parameters receive gradients
            self.opt.attach()
        self.opt.zero_grad()
        if next_batch is not None:
            self.prefetch(next_batch)
        L, _ = self._fwd_bwd(b)
        self.opt.allreduce()
        self.opt.step(clip_grad=self.clip_grad)       # per-parameter clip folded into the AdamW kernel's gradient read
        return L
