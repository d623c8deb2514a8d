"""One TAN training step on one GPU (one process per GPU): forward -> MIL-NCE / alignability loss ->
backward -> gradient all-reduce -> fused AdamW, i.e. the body of the reference's train() loop
(train/main.py:36-157: autocast forward :75-81, get_loss :83-90, backward :106, optimizer.step :111) with
the reference's optimiser policy (train/main.py:350-376,418: AdamW, no weight decay for `.ln_` / `.bias`).

B200 design:
  * parameters, gradients and AdamW moments live in flat fp32 buffers (model/params.py); the optimizer is two
    launches of exg_adamw (no-decay segment, decay segment) which also refresh the bf16 shadow weights;
  * the data-parallel replacement of DistributedDataParallel (train/main.py:413) is an NCCL all-reduce of the flat
    gradient buffer (sum; the 1/world_size of DDP's mean is folded into the AdamW kernel), issued in a few
    parameter-aligned chunks with the AdamW launch of each chunk pipelined behind its arrival, so the optimizer runs
    under the wire time of the following chunk (`comm_chunks`, `comm_dtype`);
    `cross_gpu_negatives=True` additionally all-gathers the text features so the NCE runs over the global batch
    (train/xneg.py) -- the step is then numerically the single-process step on the concatenated batch;
  * host batches land in staging buffers of the host's dtype (optionally on a copy stream while the previous step computes:
    `step(batch, next_batch=...)` / `prefetch()`), and one cast kernel per feature tensor moves them into the
    compute-dtype buffers the step reads, so the forward has no input cast and the PCIe copy overlaps compute;
  * the video-encoder and joint-encoder branches are independent given the shared input projections
    (tan_model.py:146-203), so the joint branch runs on a stream of its own, forward and backward (autograd replays
    every node on its forward stream): its latency-bound kernels fill the SMs the other branch leaves idle
    (7.98 -> 7.32 ms per step); both branches are captured into the same CUDA graph as a fork / join;
  * the device part of the step (everything between the H2D copies and the loss read-back) is static-shape
    and free of host synchronisation, so it is captured once into a CUDA graph and replayed; the random
    positional offsets of the reference (tan_model.py:157,189) are drawn on the host every step and fed to
    the graph through tiny device tensors, so replays keep the reference's augmentation semantics.
"""
import numpy as np
import torch

from .. import ops
from . import loss as exg_loss


class TanTrainStep:
    def __init__(self, model, args, B, T, N, d_video, d_text, lr=1e-4, wd=1e-5, betas=(0.9, 0.999), eps=1e-8,
                 device="cuda", use_graph=True, process_group=None, cross_gpu_negatives=False,
                 overlap_allreduce=False, bucket_mb=25.0, comm_sms=0, branch_streams=True, early_adamw=False,
                 clip_grad=0.0, comm_chunks=1, comm_dtype=None, host_dtype=torch.float32, shard_optimizer=False):
        if ops.get().name == "cuda" and not (torch.cuda.is_available() and torch.device(device).type == "cuda"):
            raise RuntimeError("exoground_b200 training needs a CUDA device (no CPU fallback)")
        self._on_cuda = torch.device(device).type == "cuda"      # False only under the tests' op-spec backend
        # 'cotrain' (train/main.py:82-89,116-117): an online / EMA-target pair.  The target's no-grad forward supplies
        # the `ema-` logits the self-labelling branch reads, and exg_ema follows the optimizer every step.
        self.twin = None
        if hasattr(model, "online") and hasattr(model, "target"):
            self.twin = model.to(device)
            model = model.online
        if getattr(args, "model", "init") == "cotrain" and self.twin is None:
            raise ValueError("args.model == 'cotrain' needs a TwinTemporalAligner (online + EMA target)")
        self.clip_grad = float(clip_grad)              # utils/train_utils.py:3-13 per-parameter clipping (0 = off)
        self.model, self.args = model.to(device), args
        self.B, self.T, self.N = B, T, N
        self.lr, self.wd, self.betas, self.eps = lr, wd, betas, eps
        self.device = torch.device(device)
        self.use_graph = use_graph and torch.device(device).type == "cuda"
        self.pg = process_group
        self.world = torch.distributed.get_world_size(process_group) if process_group is not None else 1
        # Mode B (train/xneg.py): negatives from every rank's sentences; the loss is then the GLOBAL mean, so
        # parameter gradients are summed over ranks instead of averaged
        self.xneg = None
        if cross_gpu_negatives and self.world > 1:
            from .xneg import Group
            self.xneg = Group(process_group)
            model.xneg = self.xneg
        # optional: gradient all-reduce overlapped with backward.  Every finished transformer block reports its
        # (contiguous) range of the flat gradient buffer; ranges are merged into ~bucket_mb buckets and all-reduced
        # asynchronously while the blocks below are still being differentiated.  Off by default: measured on 2 x B200
        # it hides nothing (9.40 vs 9.35 ms / step) because NCCL's CTAs cannot share an SM with the one-CTA-per-SM
        # persistent GEMMs (shared memory), so every GEMM launched while a bucket is in flight runs a second wave
        # (profiles/allreduce_overlap_r01.txt); it needs a work-stealing tile scheduler in gemm_tc2 to pay off
        self.overlap = bool(overlap_allreduce) and self.world > 1
        self.bucket_elems = int(bucket_mb * (1 << 20) / 4)
        # SMs left to NCCL's CTAs while buckets are in flight: the persistent GEMMs launched in that window use
        # (SM count - comm_sms) CTAs instead of queueing a second wave behind the SMs NCCL holds
        self.comm_sms = int(comm_sms)
        self._pending, self._works = [], []
        # multi-GPU path: after backward the flat gradient buffer is all-reduced on NCCL's stream in parameter-aligned
        # pieces (always cut at the no-decay / decay boundary; `comm_chunks` > 1 cuts further) and the AdamW launch of
        # piece i is issued as soon as piece i has arrived, i.e. under the all-reduce of piece i+1.  Measured
        # (profiles/scale_r02k_summary.txt, scale_r02r_summary.txt): more pieces do NOT pay - at 2 GPUs 1 / 4 / 8 chunks
        # are alike (the 2-rank all-reduce is itself an HBM stream), at 8 GPUs 4 chunks are SLOWER than 1 (7.92 vs
        # 7.63 ms: four 46 MB all-reduces ride lower on NCCL's latency / bandwidth curve and share HBM with AdamW) -
        # so the default is 1.  comm_dtype=torch.bfloat16 sends bf16 gradients (half the wire bytes; DDP's
        # bf16_compress_hook semantics: cast, sum in bf16 on the wire, cast back; -0.15 ms at 2 GPUs, -0.25 ms at 8),
        # the default sends the fp32 buffer like DDP does.
        # shard_optimizer=True (world > 1, no EMA twin): reduce-scatter -> AdamW on the owned 1/W of the flat buffers ->
        # all-gather of the updated bf16 shadow weights (train/flat.py::sharded_reduce_and_update).  The fp32 master
        # weights / moments of a rank are then current for its shard only until sync_full() (state_dict() calls it).
        self.shard_optimizer = bool(shard_optimizer) and self.world > 1 and self.twin is None
        self._shard_plan = None
        self.comm_chunks = max(1, int(comm_chunks))
        self.comm_dtype = comm_dtype if comm_dtype in (torch.bfloat16,) else None
        self._gcomm = None
        self._chunk_plan = None
        # the joint encoder branch on a stream of its own (model/tan_model.py::forward), forward and backward
        self.branch_stream = None
        if branch_streams and self._on_cuda and self.xneg is None:
            self.branch_stream = torch.cuda.Stream(device=self.device)
            model._branch_stream = self.branch_stream
            # parameters shared by both branches (ln_position_init, the position tables, binary_head) accumulate
            # their gradient from two streams on purpose
            quiet = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
            if quiet is not None:
                quiet(False)
        # optional, single process: the AdamW of a transformer stack is launched from inside backward as soon as the
        # stack's last block has been differentiated (on that stack's stream); lr and the bias corrections come from
        # a 3-float device buffer refreshed before every step, which keeps those launches CUDA-graph capturable
        # (exg_adamw_dev).  Off by default: measured 7.31 vs 7.34 ms per step, within run-to-run noise - only the
        # shorter (video) branch's update really overlaps, the joint branch's update sits on the critical path.
        self.early_adamw = bool(early_adamw) and self.world == 1 and self.clip_grad == 0
        self._early_on = False                 # armed only for real steps (not for prepare()'s warm-up passes)
        self._early_ranges = {}                # stack prefix -> (lo, hi, blocks) in the flat buffers
        self._early_seen, self._early_done = {}, []
        self._early_static, self._skip = [], []
        self._small_h2d = None
        self.hyper_host = torch.zeros(3, dtype=torch.float32)
        if self._on_cuda:
            self.hyper_host = self.hyper_host.pin_memory()
        self.hyper_dev = torch.zeros(3, dtype=torch.float32, device=self.device)
        self.step_count = 0
        self.graph = None
        dev = self.device
        # staging buffers (targets of the per-step H2D copies, possibly written by the copy stream while a step runs)
        # host_dtype: dtype of the feature tensors the host hands over (fp32 = what the reference's loaders produce;
        # torch.bfloat16 for feature stores kept in half precision: half the pinned pool and half the PCIe bytes)
        self.host_dtype = host_dtype
        self.video_s = torch.zeros(B, T, d_video, device=dev, dtype=host_dtype)
        self.text_s = torch.zeros(B, N, d_text, device=dev, dtype=host_dtype)
        self.vmask_s = torch.zeros(B, T, dtype=torch.bool, device=dev)
        self.tmask_s = torch.zeros(B, N, dtype=torch.bool, device=dev)
        self.tgt_s = torch.zeros(B, N, T, dtype=torch.bool, device=dev)
        self.abs_pos_s = torch.zeros(B, N, 2, device=dev)
        # static inputs of the (captured) device step; features already in the compute dtype
        Tc = model.compute_dtype
        self.video = torch.zeros(B, T, d_video, device=dev, dtype=Tc)
        self.text = torch.zeros(B, N, d_text, device=dev, dtype=Tc)
        self.vmask = torch.zeros_like(self.vmask_s)
        self.tmask = torch.zeros_like(self.tmask_s)
        self.tgt = torch.zeros_like(self.tgt_s)
        self.abs_pos = torch.zeros_like(self.abs_pos_s)
        self._copy_stream = None
        self._staged_ready = self._staging_free = None
        self._staged_id = None
        n_draw = 3 if model.use_text_pos_enc else 2
        self.pos_host = torch.zeros(n_draw, dtype=torch.int64)
        if self._on_cuda:
            self.pos_host = self.pos_host.pin_memory()
        self.pos_dev = torch.zeros(n_draw, dtype=torch.int64, device=dev)
        self.loss_out = torch.zeros((), device=dev)
        self.model._pos_offsets = [self.pos_dev[i] for i in range(n_draw)]
        self._prepared = False
        self._hooked, self._reduced = False, 0
        self._gacc, self._acc_pending = None, False        # gradient accumulation (run(apply_update=False))
        self._loss_ring, self._loss_slot = None, 0         # pinned slots of step_async()

    # ------------------------------------------------------------------ host -> device
    def _stage(self, batch, non_blocking=True):
        """H2D copies of one host batch into the staging buffers, on the current stream.  Returns bytes copied."""
        n = 0
        for dst, key in ((self.video_s, "video"), (self.text_s, "text"), (self.vmask_s, "video_padding_mask"),
                         (self.tmask_s, "text_padding_mask")):
            src = batch[key]
            if tuple(src.shape) != tuple(dst.shape):
                # the captured step has static shapes; a silent broadcast (B' = 1) or a copy_ error deep inside is
                # worse than saying so.  Ragged batches: pad to the configured (B, T, N) with masked rows.
                raise ValueError(f"TanTrainStep: batch[{key!r}] has shape {tuple(src.shape)}, the step was built for "
                                 f"{tuple(dst.shape)} (B={self.B}, T={self.T}, N={self.N}); pad the batch or build "
                                 f"another step")
            if src.dtype != dst.dtype:
                raise ValueError(f"TanTrainStep: batch[{key!r}] is {src.dtype}, the step was built with "
                                 f"host_dtype={dst.dtype} (a converting copy would run on the host, inside the step)")
            dst.copy_(src, non_blocking=non_blocking)
            n += src.numel() * src.element_size()
        if "tgt" in batch:
            tgt = batch["tgt"]
        else:
            tgt = exg_loss.get_mask_from_time(batch["start"], batch["end"], self.T, self.N, device="cpu")[0]
        if tuple(tgt.shape) != tuple(self.tgt_s.shape):
            raise ValueError(f"TanTrainStep: target mask has shape {tuple(tgt.shape)}, expected "
                             f"{tuple(self.tgt_s.shape)}")
        self.tgt_s.copy_(tgt, non_blocking=non_blocking)
        n += tgt.numel()
        if batch.get("abs_text_pos") is not None:
            self.abs_pos_s.copy_(batch["abs_text_pos"], non_blocking=non_blocking)
            n += self.abs_pos_s.numel() * 4
        return n

    def _commit(self):
        """staging -> static step inputs on the current stream (feature tensors: one cast kernel each)."""
        K = ops.get()
        for src, dst in ((self.video_s, self.video), (self.text_s, self.text)):
            if dst.dtype == src.dtype:
                dst.copy_(src)
            else:
                K.cast(src, dst.dtype, out=dst)
        self.vmask.copy_(self.vmask_s)
        self.tmask.copy_(self.tmask_s)
        self.tgt.copy_(self.tgt_s)
        self.abs_pos.copy_(self.abs_pos_s)

    def load_batch(self, batch, non_blocking=True):
        """Copies one host batch (dict of synth.tan_batch / the loaders' collate dict; pinned tensors give
        asynchronous copies) into the device buffers the step reads and builds the [B,N,T] target mask.
        Returns the number of bytes copied host->device."""
        if self._staging_free is not None:
            torch.cuda.current_stream().wait_event(self._staged_ready)     # never race a prefetch in flight
        n = self._stage(batch, non_blocking)
        self._commit()
        if self._staging_free is not None:
            self._staging_free.record()
        self._staged_id = None
        return n

    def prefetch(self, batch):
        """Starts the H2D copies of a FUTURE batch on a side stream; they overlap whatever the main stream is
        running.  The next step(batch) / load_batch of the same dict object picks the staged copy up."""
        if not self._on_cuda:
            self._stage(batch, False)
            self._staged_id = id(batch)
            return
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream()
            self._staged_ready = torch.cuda.Event()
            self._staging_free = torch.cuda.Event()
            self._staging_free.record()
        cs = self._copy_stream
        cs.wait_event(self._staging_free)           # the last _commit() has read the staging buffers
        with torch.cuda.stream(cs):
            self._stage(batch, True)
            self._staged_ready.record(cs)
        self._staged_id = id(batch)

    def _take_staged(self):
        if self._on_cuda:
            torch.cuda.current_stream().wait_event(self._staged_ready)
        self._commit()
        if self._on_cuda:
            self._staging_free.record()
        self._staged_id = None

    def _draw_offsets(self):
        m = self.model
        if m.random_pos_start:
            self.pos_host[0] = int(np.random.randint(0, int(self.T / 2)))
            k = 1
            if m.use_text_pos_enc:
                self.pos_host[k] = int(np.random.randint(0, int(self.N / 2)))
                k += 1
            self.pos_host[k] = int(np.random.randint(0, int(self.T / 2)))
        self.pos_dev.copy_(self.pos_host, non_blocking=True)

    # ------------------------------------------------------------------ device step
    def _fwd_bwd(self):
        m = self.model
        out = m(self.video, self.text, self.vmask, self.tmask, None)
        if self.twin is not None and getattr(self.args, "model", "init") == "cotrain":
            with torch.no_grad():          # train/main.py:82-89 (the target never draws a random offset, tan_model.py:326)
                ema = self.twin.target(self.video, self.text, self.vmask, self.tmask, None)
            out = {**out, **{f"ema-{k}": v for k, v in ema.items()}}
        L = exg_loss.loss_core(self.tgt, self.tmask, out, self.args,
                               self.abs_pos if getattr(self.args, "use_abs_text_pos", False) else None,
                               self.B, self.T, self.N, video_padding_mask=self.vmask, xneg=self.xneg)
        L["loss"].backward()
        if self.branch_stream is not None:     # every gradient written on the branch stream is final past this point
            torch.cuda.current_stream().wait_stream(self.branch_stream)
        self.loss_out.copy_(L["loss-global" if self.xneg is not None else "loss"].detach())
        if self._hooked:
            self._finish_allreduce()

    # ------------------------------------------------------------------ overlapped gradient all-reduce
    def _on_block_done(self, prefix):
        if self.early_adamw:
            return self._early_block_done(prefix)
        r = self.model.flat_params().layer_ranges.get(prefix)
        if r is None:
            return
        self._pending.append(r)
        if sum(hi - lo for lo, hi in self._pending) >= self.bucket_elems:
            self._flush_pending()

    def _flush_pending(self):
        g = self.model.flat_params().grad
        merged = []
        for lo, hi in sorted(self._pending):
            if merged and merged[-1][1] == lo:
                merged[-1][1] = hi
            else:
                merged.append([lo, hi])
        if merged and self.comm_sms > 0 and not self._works:
            ops.get().reserve_sms(self.comm_sms)
        for lo, hi in merged:
            self._works.append(torch.distributed.all_reduce(g[lo:hi], group=self.pg, async_op=True))
            self._reduced += hi - lo
        self._pending = []

    def _finish_allreduce(self):
        fp = self.model.flat_params()
        self._flush_pending()
        if self._reduced != fp.n_used - fp.n_bucketed:
            raise RuntimeError("overlapped all-reduce: a transformer block did not report its gradients")
        if fp.n_bucketed > 0:
            self._works.append(torch.distributed.all_reduce(fp.grad[:fp.n_bucketed], group=self.pg, async_op=True))
        for w in self._works:
            w.wait()
        if self.comm_sms > 0:
            ops.get().reserve_sms(0)
        self._works, self._reduced = [], 0

    # ------------------------------------------------------------------ AdamW launched from inside backward
    def _early_block_done(self, prefix):
        if not self._early_on:
            return
        stack = prefix[:prefix.index("resblocks.")]
        rng = self._early_ranges.get(stack)
        if rng is None:
            return
        self._early_seen[stack] = self._early_seen.get(stack, 0) + 1
        if self._early_seen[stack] != rng[2]:
            return
        fp = self.model.flat_params()
        lo, hi = rng[0], rng[1]
        b1, b2 = self.betas
        shadow = fp.shadow if self.model.compute_dtype == torch.bfloat16 else None
        ops.get().adamw_dev(fp.flat[lo:hi], fp.grad[lo:hi], self.m1[lo:hi], self.m2[lo:hi],
                            shadow[lo:hi] if shadow is not None else None, self.hyper_dev, b1, b2, self.eps, self.wd)
        self._early_done.append((lo, hi))

    def _set_hyper(self, step):
        b1, b2 = self.betas
        self.hyper_host[0] = self.lr
        self.hyper_host[1] = 1.0 - b1 ** step
        self.hyper_host[2] = (1.0 - b2 ** step) ** 0.5
        self.hyper_dev.copy_(self.hyper_host, non_blocking=True)

    def _optimizer(self):
        fp = self.model.flat_params()
        K = ops.get()
        self.step_count += 1
        b1, b2 = self.betas
        shadow = fp.shadow if self.model.compute_dtype == torch.bfloat16 else None
        segs = [(0, fp.n_no_decay, 0.0)]
        cur = fp.n_no_decay                      # decay segment minus the ranges already updated inside backward
        skip, self._skip = sorted(self._skip), []
        for lo, hi in skip:
            if lo > cur:
                segs.append((cur, lo, self.wd))
            cur = max(cur, hi)
        if fp.n_used > cur:
            segs.append((cur, fp.n_used, self.wd))
        gscale = 1.0 if self.xneg is not None else 1.0 / self.world
        if self.clip_grad > 0:                         # train/main.py:108-110, fused: norms in one launch, factor in AdamW
            K.grad_sumsq(fp.grad[:fp.n_used], self.blk2seg, len(self.seg_names), out=self.seg_sumsq)
        for lo, hi, wd in segs:
            if hi > lo:
                sh = shadow[lo:hi] if shadow is not None else None
                if self.clip_grad > 0:
                    K.adamw_clip(fp.flat[lo:hi], fp.grad[lo:hi], self.m1[lo:hi], self.m2[lo:hi], sh, self.lr, b1, b2,
                                 self.eps, wd, self.step_count, self.blk2seg[lo // 64:], self.seg_sumsq,
                                 self.clip_grad, grad_scale=gscale)
                else:
                    K.adamw(fp.flat[lo:hi], fp.grad[lo:hi], self.m1[lo:hi], self.m2[lo:hi], sh, self.lr, b1, b2,
                            self.eps, wd, self.step_count, grad_scale=gscale)
        if self.twin is not None:                      # train/main.py:116-117
            self.twin._momentum_update()

    # ------------------------------------------------------------------ checkpointing (train/main.py:136-146,469)
    def sync_full(self):
        """shard_optimizer: completes the fp32 master weights and the AdamW moments on every rank (collective; call on
        all ranks before model.state_dict() / state_dict()).  No-op otherwise."""
        if self._shard_plan is not None:
            from .flat import gather_full
            gather_full(self._shard_plan, self.pg, self.model.flat_params().flat, self.m1, self.m2)

    def state_dict(self):
        """Optimizer state in torch.optim.AdamW's layout over the reference's two parameter groups: what the
        reference stores under 'optimizer' in its runtime / epoch checkpoints, so training on the fused path can be
        checkpointed and resumed like (and interchangeably with) the reference loop."""
        from .flat import optim_state_dict
        if not self._prepared:
            self.prepare()
        self.sync_full()
        return optim_state_dict(self.model, self.m1, self.m2, self.step_count, self.lr, self.wd, self.betas, self.eps)

    def load_state_dict(self, sd):
        """Restores AdamW moments / step count (and lr, weight decay) from state_dict() or from a checkpointed
        torch.optim.AdamW over the reference's parameter groups.  Load the MODEL's state_dict first."""
        from .flat import load_optim_state_dict
        if not self._prepared:
            self.prepare()
        self.step_count, self.lr, self.wd = load_optim_state_dict(self.model, self.m1, self.m2, sd)
        if self.model.compute_dtype == torch.bfloat16:
            self.model.flat_params().refresh_shadow()      # the graph replays read the shadow, not the fp32 weights

    def _reduce_and_update(self):
        """Chunked gradient all-reduce with the optimizer pipelined behind it (see __init__; train/flat.py)."""
        from .flat import plan_chunks, reduce_and_update, ShardPlan, sharded_reduce_and_update
        fp = self.model.flat_params()
        if self.shard_optimizer:
            if self._shard_plan is None:
                self._shard_plan = ShardPlan(fp, self.world, torch.distributed.get_rank(self.pg), self.wd)
                if self.comm_dtype is not None:
                    self._gcomm = torch.empty(self._shard_plan.padded, dtype=self.comm_dtype, device=fp.flat.device)
            self.step_count += 1
            sharded_reduce_and_update(fp, self._shard_plan, self.pg, self.m1, self.m2, self.lr, self.betas, self.eps,
                                      self.step_count, 1.0 if self.xneg is not None else 1.0 / self.world,
                                      self.clip_grad, self.blk2seg, len(self.seg_names), self.seg_sumsq, self._gcomm,
                                      shadow=self.model.compute_dtype == torch.bfloat16)
            return
        if self._chunk_plan is None:
            self._chunk_plan = plan_chunks(fp, self.comm_chunks, self.wd)
            if self.comm_dtype is not None:
                self._gcomm = torch.empty(fp.n_used, dtype=self.comm_dtype, device=fp.flat.device)
        self.step_count += 1
        reduce_and_update(fp, self._chunk_plan, self.pg, self.m1, self.m2, self.lr, self.betas, self.eps,
                          self.step_count, 1.0 if self.xneg is not None else 1.0 / self.world, self.clip_grad,
                          self.blk2seg, len(self.seg_names), self.seg_sumsq, self._gcomm,
                          shadow=self.model.compute_dtype == torch.bfloat16)
        if self.twin is not None:                      # train/main.py:116-117
            self.twin._momentum_update()

    def prepare(self):
        """Discovery pass (which parameters get gradients), flat gradient / moment buffers, kernel warm-up and,
        if enabled, CUDA-graph capture of forward+loss+backward.  Uses whatever is in the static buffers."""
        m = self.model
        m.train()
        for p in m.parameters():
            p.grad = None
        self._draw_offsets()
        self._fwd_bwd()
        unused = [n for n, p in m.named_parameters() if p.grad is None]
        fp = m.flat_params()
        fp.set_unused(unused)
        fp = m.flat_params()
        m._shadow_stale = True
        if self.twin is not None:
            # the EMA target follows the online model's flat layout (exg_ema runs buffer against buffer); settle it
            # BEFORE the capture below, so no replay ever reads a re-laid-out (freed) target buffer
            self.twin.target._flat.set_unused(fp.unused)
            self.twin.target.flat_params()
            self.twin.target._shadow_stale = True
        g = fp.grad_buffer()
        for n, p in m.named_parameters():
            off, numel, shape = fp.index[n]
            p.grad = g[off:off + numel].view(shape) if n not in fp.unused else None
        for mod in m.modules():           # stack weight gradients go straight into the flat buffer (engine sinks)
            mod._grad_sink = True
            if self.overlap and hasattr(mod, "resblocks"):
                mod._grad_ready_hook = self._on_block_done
        self._hooked = self.overlap and len(fp.layer_ranges) > 0
        self._early_static = []
        if self.early_adamw:
            stacks = {}
            for pre, (lo, hi) in fp.layer_ranges.items():
                st = pre[:pre.index("resblocks.")]
                a = stacks.setdefault(st, [lo, hi, 0])
                a[0], a[1], a[2] = min(a[0], lo), max(a[1], hi), a[2] + 1
            for mod in m.modules():
                if hasattr(mod, "resblocks") and hasattr(mod, "run") and type(mod).__name__ == "TemporalEncoder":
                    mod._grad_ready_hook = self._on_block_done
            # only stacks whose blocks form one contiguous range and report through the encoder engine
            enc = {n + "." for n, mod in m.named_modules() if type(mod).__name__ == "TemporalEncoder"}
            for st, (lo, hi, nb) in stacks.items():
                covered = sum(h - l for p_, (l, h) in fp.layer_ranges.items() if p_.startswith(st))
                if st in enc and covered == hi - lo:
                    self._early_ranges[st] = (lo, hi, nb)
            self._early_static = sorted((lo, hi) for lo, hi, _ in self._early_ranges.values())
        self.m1 = torch.zeros_like(fp.flat)
        self.m2 = torch.zeros_like(fp.flat)
        from .flat import build_blk2seg
        self.blk2seg, self.seg_names = build_blk2seg(fp)
        self.seg_sumsq = torch.zeros(len(self.seg_names), dtype=torch.float32, device=fp.flat.device)
        if m.compute_dtype == torch.bfloat16:
            fp.refresh_shadow()
            m._shadow_stale = False
            # from here on the shadow is kept fresh by the AdamW kernel itself
            m._shadow_managed = True
        if self._on_cuda:
            torch.cuda.synchronize()
        if self.use_graph:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):
                    g.zero_()
                    self._fwd_bwd()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            self._early_on = self.early_adamw
            self._early_seen, self._early_done = {}, []
            with torch.cuda.graph(self.graph):
                g.zero_()
                self._fwd_bwd()
            self._early_on = False
            self._early_static = sorted(self._early_done)     # what the captured backward really launched
            torch.cuda.synchronize()
        g.zero_()
        self._prepared = True

    def close(self):
        """Release the captured graph.  Call before torch.distributed.destroy_process_group(): a CUDA graph that
        captured NCCL kernels (cross_gpu_negatives) must not outlive its communicator."""
        if self._on_cuda:
            torch.cuda.synchronize()
        self.graph = None
        if self._on_cuda:
            torch.cuda.synchronize()

    def run(self, apply_update=True):
        """Device part of one step on the current static buffers.  Returns the 0-d loss tensor (device).
        apply_update=False: gradient accumulation (train/main.py:106-113, `backprop_freq`): forward + loss + backward
        only; this micro-batch's gradient is ADDED to an accumulator (the reference neither divides the loss nor the
        gradient by backprop_freq) and the next call with apply_update=True clips / all-reduces / applies the sum."""
        if not self._prepared:
            self.prepare()
        if (self.overlap or self.early_adamw) and (not apply_update or self._acc_pending):
            raise ValueError("gradient accumulation needs the post-backward gradient exchange (no overlap_allreduce / "
                             "early_adamw): their collectives / updates run inside backward")
        # the tiny pinned buffers below are rewritten every step: wait until the previous step's asynchronous copies
        # have executed (i.e. the step before it has finished), so a host that runs ahead cannot overwrite them early
        if self._small_h2d is not None:
            self._small_h2d.synchronize()
        self._draw_offsets()
        if self.early_adamw:
            self._set_hyper(self.step_count + 1)
        if self._on_cuda:
            if self._small_h2d is None:
                self._small_h2d = torch.cuda.Event()
            self._small_h2d.record()
        if self.graph is not None:
            self.graph.replay()
            self._skip = list(self._early_static)          # the replay has updated these ranges
        else:
            self.model.flat_params().grad.zero_()
            self._early_on = self.early_adamw
            self._early_seen, self._early_done = {}, []
            self._fwd_bwd()
            self._early_on = False
            self._skip = list(self._early_done)
        if not apply_update or self._acc_pending:
            g = self.model.flat_params().grad
            if self._gacc is None:
                self._gacc = torch.zeros_like(g)
            self._gacc.add_(g)
            self._acc_pending = True
            if not apply_update:
                return self.loss_out
            g.copy_(self._gacc)                    # the sum over the micro-batches is what gets clipped and applied
            self._gacc.zero_()
            self._acc_pending = False
        if self.world > 1 and not self._hooked:
            self._reduce_and_update()
        else:
            self._optimizer()
        return self.loss_out

    class _PendingLoss:
        """Handle of step_async(): the loss of that step, read back without stalling the stream."""

        def __init__(self, host, event):
            self._host, self._event = host, event

        def result(self):
            if self._event is not None:
                self._event.synchronize()
            return float(self._host.item())

    def step_async(self, batch, next_batch=None, apply_update=True):
        """step() without the host stall: the loss goes to a pinned host slot by an asynchronous copy and the call
        returns a handle (`.result()` -> float).  A loop that reads the PREVIOUS step's handle after launching the next
        step keeps the GPU busy through the host's launch work (the reference logs `loss.item()` every iteration,
        train/main.py:120-126; reading it one step late changes nothing else).  Two slots: a handle must be read before
        the second call after the one that produced it."""
        if self._staged_id is not None and self._staged_id == id(batch):
            self._take_staged()
        else:
            self.load_batch(batch)
        loss = self.run(apply_update)
        if next_batch is not None:
            self.prefetch(next_batch)
        if not self._on_cuda:
            return TanTrainStep._PendingLoss(loss.detach().clone().cpu(), None)
        if self._loss_ring is None:
            self._loss_ring = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
            self._loss_slot = 0
        host = self._loss_ring[self._loss_slot]
        self._loss_slot ^= 1
        host.copy_(loss, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        return TanTrainStep._PendingLoss(host, ev)

    def step(self, batch, next_batch=None, apply_update=True):
        """The user-facing call: host batch in, python float loss out (H2D copies and the loss read-back
        included).  `next_batch` (optional) is staged on the copy stream while this step computes, so a loop
        `for b, nb in zip(batches, batches[1:]): step(b, next_batch=nb)` hides the PCIe copy behind the step."""
        if self._staged_id is not None and self._staged_id == id(batch):
            self._take_staged()
        else:
            self.load_batch(batch)
        loss = self.run(apply_update)
        if next_batch is not None:
            self.prefetch(next_batch)
        return float(loss.item())
