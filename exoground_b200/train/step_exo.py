"""One ExoGround / keystep-grounding training step on one GPU (one process per GPU): the body of the reference's
train() loop (train/main_egoexo4d_distributed.py:105-222: forward :145-153, get_loss :165-168, backward :175,
clip + optimizer step :176-181) with DistributedDataParallel (:611-614) replaced by a chunked NCCL all-reduce of the
flat gradient buffer with the fused clip + AdamW of each chunk pipelined behind its arrival (train/flat.py::
reduce_and_update) instead of torch.optim.AdamW.  The grounding / view-invariant losses
have no cross-sample terms, so the data-parallel mode is DDP-equivalent: local loss, mean gradient.

Device step.  For the `grounding` / `joint` models the whole device part - forward, the interval-regression + IoU loss
(one kernel), the cross-view InfoNCE (one kernel), backward - has static shapes and no host synchronisation: the
statistics the reference reads with one `.item()` per value (loss_egoexo4d.py:121-122,177-186) are written to one small
device buffer and read back ONCE after the step, and the reference's loss dictionary (same keys) is built from that
read-back on the host.  That makes the step capturable: `use_graph=True` (default on CUDA) replays forward + loss +
backward as one CUDA graph, which removes the ~1000 eager launches per step that made 8 ranks on one host
launch-bound (4.9x at 8 GPUs in round 1).  The reference's random positional offsets are drawn on the host every step
and fed to the graph through device scalars.  `args.test`, the `view_invariant` model and shapes that change between
batches take the eager path through the drop-in get_loss."""
import numpy as np
import torch

from .. import ops
from . import loss_egoexo4d as exo_loss
from .flat import FlatAdamW

_DEVICE_KEYS = ("video_features", "narration_features", "video_padding_mask", "narration_padding_mask",
                "ego_video_features", "view_rank_label", "view_rank_neg_label", "same_view_neg_idxs",
                "starts", "ends", "mean", "duration", "view_available_mask")


class ExoTrainStep:
    def __init__(self, model, args, lr=1e-4, wd=1e-5, betas=(0.9, 0.999), eps=1e-8, clip_grad=0.0, device="cuda",
                 process_group=None, branch_streams=True, use_graph=True, shard_optimizer=False):
        if ops.get().name == "cuda" and not (torch.cuda.is_available() and torch.device(device).type == "cuda"):
            raise RuntimeError("exoground_b200 training needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        self.model, self.args, self.clip_grad = model.to(device), args, clip_grad
        self.opt = FlatAdamW(self.model, lr, wd, betas, eps, process_group, shard=shard_optimizer)
        self.h2d_bytes = 0
        self._copy_stream, self._staged = None, None
        self._on_cuda = self.device.type == "cuda" and ops.get().name == "cuda"
        # the device-loss path (one read-back per step; capturable)
        distill = bool(getattr(args, "use_distill_nce_loss", False))
        self._joint_distill = args.model == "joint" and distill
        self.device_loss = (args.model in ("grounding", "joint") and not getattr(args, "test", False)
                            and not getattr(args, "test_egovlp", False))
        self.use_graph = bool(use_graph) and self._on_cuda and self.device_loss
        self.graph = None
        self._static = None
        self._shapes = None
        # text unimodal encoder beside the video unimodal encoder (model/exo_ground_model.py::_unimodal_pair)
        self.branch_stream = None
        if branch_streams and self._on_cuda:
            self.branch_stream = torch.cuda.Stream(device=self.device)
            self.model._branch_stream = self.branch_stream
            quiet = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
            if quiet is not None:
                quiet(False)
        # device-side positional offsets for graph replays (the model consumes them inside forward() only)
        from ..model.exo_ground_model import ExoGroundingTransformer
        self._n_draw = 0
        if self.use_graph and getattr(model, "random_pos_start", 0):
            if type(model) is not ExoGroundingTransformer:
                self.use_graph = False         # sub-classes draw on the host inside their own forward: stay eager
            else:
                self._n_draw = 2 + (1 if model.use_text_pos_enc else 0)
        if self._n_draw:
            self.pos_host = torch.zeros(self._n_draw, dtype=torch.int64).pin_memory()
            self.pos_dev = torch.zeros(self._n_draw, dtype=torch.int64, device=self.device)
            self._pos_event = None

    # ------------------------------------------------------------------ host -> device
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

    # ------------------------------------------------------------------ device step
    def _model_out(self, b):
        m, args = self.model, self.args
        ego = b.get("ego_video_features") if getattr(args, "use_distill_nce_loss", False) else None
        kw = {}
        if ego is not None:
            kw["egocentric_video_embed"] = ego[:, 0] if ego.dim() == 4 else ego
        return m(b["video_features"], b["narration_features"], b["video_padding_mask"].bool(),
                 b["narration_padding_mask"].bool(), **kw)

    def _fwd_bwd(self, b):
        """Eager path through the drop-in get_loss (its host read-backs included)."""
        out = self._model_out(b)
        L, iou = exo_loss.get_loss(b, out, b["narration_padding_mask"], self.args)
        L["loss"].backward()
        if self.branch_stream is not None:
            torch.cuda.current_stream().wait_stream(self.branch_stream)
        return L, iou

    def _device_fwd_bwd(self, b):
        """forward + device-side losses + backward, no host synchronisation.  Returns (loss, stats, per-second
        monitors [4, B*T] or None) as device tensors."""
        args = self.args
        out = self._model_out(b)
        tmask = b["narration_padding_mask"]
        if args.use_center_duration:
            gt_a, gt_b = b["mean"], b["duration"]
        else:
            gt_a, gt_b = b["starts"], b["ends"]
        loss, stats, _ = exo_loss.ground_loss_core(out["interval_preds"], gt_a, gt_b, tmask, args.use_center_duration,
                                                   args.iou_loss_eps, list(args.iou_thresholds))
        per_sec = None
        if self._joint_distill:
            svn = getattr(args, "same_view_negative", False)
            nce, l1, pc, nc = exo_loss._xview(out["high_dim_features"], b["ego_video_features"], b["view_rank_label"],
                                              b["view_rank_neg_label"], b["same_view_neg_idxs"] if svn else None,
                                              b["video_features"] if svn else None,
                                              getattr(args, "only_same_view_negative", False), exo_loss.INV_TEMP)
            per_sec = torch.stack((l1, pc, nc, nce.detach()), dim=0).reshape(4, -1)
            loss = nce.mean() + loss
        loss.backward()
        if self.branch_stream is not None:
            torch.cuda.current_stream().wait_stream(self.branch_stream)
        return loss.detach(), stats, per_sec

    def _loss_dict(self, loss, stats, per_sec, batch):
        """The reference's loss dictionary (loss_egoexo4d.py:102-139,173-198,349-363) from ONE read-back."""
        args = self.args
        stats_h = stats.detach().float().cpu()
        d = {}
        if self._joint_distill:
            host = per_sec.detach().double().cpu().numpy()
            means = tuple(torch.tensor(float(h.mean()), dtype=torch.float32) for h in host)
            d.update(exo_loss.view_invariant_dict(host, batch["metadata"]["per_second_views"], means))
        d.update(exo_loss.grounding_dict(stats_h, stats_h.tolist(), args.use_center_duration, list(args.iou_thresholds)))
        d["loss"] = loss.detach().float().cpu() if loss.is_cuda else loss.detach()
        return d

    def _draw_offsets(self):
        m = self.model
        if not self._n_draw:
            return
        if self._pos_event is not None:
            self._pos_event.synchronize()          # the previous step's copy of the pinned scalars has executed
        T = self._shapes["video_features"][1] // m.num_max_views
        N = self._shapes["narration_features"][1]
        k = 0
        if m.use_text_pos_enc:                     # draw order of the reference forward: text, video unimodal, joint
            self.pos_host[k] = int(np.random.randint(0, int(N / 2))); k += 1
        self.pos_host[k] = int(np.random.randint(0, int(T / 2))); k += 1
        self.pos_host[k] = int(np.random.randint(0, int(T / 2)))
        self.pos_dev.copy_(self.pos_host, non_blocking=True)
        if self._pos_event is None:
            self._pos_event = torch.cuda.Event()
        self._pos_event.record()

    def _keys_used(self, b):
        return [k for k in _DEVICE_KEYS if torch.is_tensor(b.get(k))]

    def _capture(self, b):
        """Static input buffers shaped like the first batch, warm-up, capture of forward + loss + backward."""
        self._shapes = {k: tuple(b[k].shape) for k in self._keys_used(b)}
        self._static = {k: torch.empty_like(b[k]) for k in self._shapes}
        for k, v in self._static.items():
            v.copy_(b[k])
        if self._n_draw:
            # (warm-up and capture run on whatever pos_dev holds - zeros; the real offsets are drawn per step, in the
            # reference's order, so an eager run with the same numpy seed sees the same sequence)
            self.model._pos_offsets = [self.pos_dev[i] for i in range(self._n_draw)]
        sb = dict(b)
        sb.update(self._static)
        self._static_batch = sb
        g = self.opt.flat_grad()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                self.opt.zero_grad()
                self._device_fwd_bwd(sb)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.opt.zero_grad()
            loss, stats, per_sec = self._device_fwd_bwd(sb)
            self._out = (loss.clone(), stats.clone(), per_sec.clone() if per_sec is not None else None)
        torch.cuda.synchronize()
        g.zero_()

    def close(self):
        if self._on_cuda:
            torch.cuda.synchronize()
        self.graph = None

    def step(self, batch, next_batch=None, apply_update=True):
        """Host batch (the collate dict of data/loader_egoexo4d.py:567-598) in, loss dict out.  `next_batch`
        (optional) is copied to the device on the copy stream while this step computes.  apply_update=False:
        gradient accumulation (`backprop_freq`, main_egoexo4d_distributed.py:176) - forward + loss + backward only."""
        self.model.train()
        b = self._take(batch)
        if not self.opt.attached:
            for p in self.model.parameters():
                p.grad = None
            self._fwd_bwd(b)                    # discovery pass: which parameters receive gradients
            self.opt.attach()
        if not self.device_loss:
            self.opt.zero_grad()
            if next_batch is not None:
                self.prefetch(next_batch)
            L, _ = self._fwd_bwd(b)
        elif self.use_graph:
            if self.graph is None:
                self._capture(b)
            shapes = {k: tuple(b[k].shape) for k in self._keys_used(b)}
            if shapes != self._shapes:
                raise ValueError(f"ExoTrainStep: the captured step was built for {self._shapes}, got {shapes}; pad the "
                                 f"batch or construct the step with use_graph=False")
            for k, v in self._static.items():
                v.copy_(b[k], non_blocking=True)
            self._draw_offsets()
            if next_batch is not None:
                self.prefetch(next_batch)
            self.graph.replay()
            if self.opt.accumulate(apply_update):
                self.opt.reduce_and_step(clip_grad=self.clip_grad)
            return self._loss_dict(self._out[0], self._out[1], self._out[2], batch)
        else:
            self.opt.zero_grad()
            if next_batch is not None:
                self.prefetch(next_batch)
            loss, stats, per_sec = self._device_fwd_bwd(b)
            if self.opt.accumulate(apply_update):
                self.opt.reduce_and_step(clip_grad=self.clip_grad)
            return self._loss_dict(loss, stats, per_sec, batch)
        if self.opt.accumulate(apply_update):
            self.opt.reduce_and_step(clip_grad=self.clip_grad)   # per-parameter clip folded into AdamW's gradient read
        return L
