"""Optimizer side of a data-parallel step over the flat parameter / gradient buffers of an ExgModule
(model/params.py): the replacement of DistributedDataParallel + torch.optim.AdamW + GradScaler of the reference
trainers (train/main_egoexo4d_distributed.py:544-570 two-group policy — no weight decay for `.ln_` / `.bias` —,
:721 AdamW, :176-181 clip + step; utils/train_utils.py:3-13 per-parameter clipping).

  * `attach()` after one discovery backward: parameters that received no gradient are moved out of the optimised
    range, every `.grad` becomes a view into ONE flat fp32 buffer, the transformer stacks write their weight
    gradients straight into it (engine "sinks"), AdamW moments are flat too;
  * `allreduce()`: ONE NCCL all-reduce of the flat gradient (sum; DDP's 1/world is folded into the AdamW kernel);
  * `clip_per_param()`: the reference's per-parameter norm clipping without its one `.item()` per parameter;
  * `step()`: two launches of the fused AdamW kernel (no-decay segment, decay segment), which also refresh the
    bf16 shadow weights the GEMMs read.
"""
import torch

from .. import ops
from ..model.params import is_no_decay, ALIGN


def build_blk2seg(fp):
    """(int32 [n_used / 64] block -> parameter table, parameter names in table order) for the optimised range of a
    FlatParams: parameters start on 64-element boundaries, the padding after a parameter belongs to it (zeros)."""
    names = [n for n, (off, numel, _) in sorted(fp.index.items(), key=lambda kv: kv[1][0]) if off < fp.n_used]
    tab = torch.zeros(fp.n_used // ALIGN, dtype=torch.int32)
    for i, n in enumerate(names):
        off, numel, _ = fp.index[n]
        tab[off // ALIGN:(off + numel + ALIGN - 1) // ALIGN] = i
    return tab.to(fp.flat.device), names


def policy_groups(model):
    """The reference's two AdamW groups in ITS order (train/main.py:350-376, main_egoexo4d_distributed.py:544-570:
    `named_parameters()` order, requires_grad only; no weight decay for `.ln_` / `.bias` names of the DDP-wrapped
    model): ([(name, param)] no-decay, [(name, param)] decay)."""
    nd, dc = [], []
    for n, p in model.named_parameters():
        if p.requires_grad:
            (nd if is_no_decay(n) else dc).append((n, p))
    return nd, dc


def optim_state_dict(model, m1, m2, step_count, lr, wd, betas, eps):
    """The flat AdamW state in torch.optim.AdamW's `state_dict()` layout for the reference's two parameter groups, so
    the reference's checkpoints (`'optimizer': optimizer.state_dict()`, train/main.py:136-146) can be written from
    and resumed into the fused path.  Parameters that never received a gradient carry no state (as in torch)."""
    nd, dc = policy_groups(model)
    opt = torch.optim.AdamW([{"params": [p for _, p in nd], "lr": lr, "weight_decay": 0.0},
                             {"params": [p for _, p in dc], "lr": lr, "weight_decay": wd}],
                            lr=lr, weight_decay=wd, betas=betas, eps=eps)
    sd = opt.state_dict()
    fp = model.flat_params()
    state = {}
    if step_count > 0:
        for i, (n, _) in enumerate(nd + dc):
            if n in fp.unused or n not in fp.index:
                continue
            off, numel, shape = fp.index[n]
            state[i] = {"step": torch.tensor(float(step_count)),
                        "exp_avg": m1[off:off + numel].view(shape).clone(),
                        "exp_avg_sq": m2[off:off + numel].view(shape).clone()}
    sd["state"] = state
    return sd


def load_optim_state_dict(model, m1, m2, sd):
    """Inverse of optim_state_dict (also accepts a state_dict produced by torch.optim.AdamW over the reference's
    groups).  Returns (step_count, lr, weight_decay of the decay group)."""
    nd, dc = policy_groups(model)
    named = nd + dc
    groups = sd["param_groups"]
    n_params = sum(len(g["params"]) for g in groups)
    if n_params != len(named):
        raise ValueError(f"optimizer state has {n_params} parameters, the model's policy groups have {len(named)}")
    fp = model.flat_params()
    m1.zero_()
    m2.zero_()
    steps = set()
    for i, st in sd["state"].items():
        n = named[int(i)][0]
        off, numel, shape = fp.index[n]
        if tuple(st["exp_avg"].shape) != tuple(shape):
            raise ValueError(f"optimizer state of {n}: shape {tuple(st['exp_avg'].shape)} != {tuple(shape)}")
        m1[off:off + numel].view(shape).copy_(st["exp_avg"])
        m2[off:off + numel].view(shape).copy_(st["exp_avg_sq"])
        steps.add(int(float(st["step"])))
    if len(steps) > 1:
        raise ValueError(f"optimizer state carries different step counts {sorted(steps)}: the fused AdamW keeps one")
    return (steps.pop() if steps else 0), float(groups[-1]["lr"]), float(groups[-1]["weight_decay"])


def plan_chunks(fp, n_chunks, wd):
    """[(lo, hi, weight decay)]: the used part of the flat buffer cut at parameter boundaries into about `n_chunks`
    equal pieces (a parameter never straddles two chunks, so its clipping norm is complete when its chunk is updated);
    the no-decay / decay boundary is always a cut."""
    starts = sorted(off for n, (off, numel, _) in fp.index.items() if n not in fp.unused and off < fp.n_used)
    want = max(1, (fp.n_used + n_chunks - 1) // max(1, n_chunks))
    cuts, last = [0], 0
    for off in starts[1:] + [fp.n_used]:
        if (off == fp.n_no_decay or off - last >= want or off == fp.n_used) and off > last:
            cuts.append(off)
            last = off
    plan = [(lo, hi, 0.0 if hi <= fp.n_no_decay else wd) for lo, hi in zip(cuts[:-1], cuts[1:])]
    assert plan and plan[0][0] == 0 and plan[-1][1] == fp.n_used
    return plan


def reduce_and_update(fp, plan, pg, m1, m2, lr, betas, eps, step, gscale, clip, blk2seg, n_seg, seg_sumsq, gcomm=None,
                      shadow=False):
    """The data-parallel replacement of DDP's bucketed all-reduce + optimizer.step() (train/main.py:413,111;
    main_egoexo4d_distributed.py:611-614,181): every chunk of `plan` is all-reduced asynchronously on NCCL's stream and
    the (clip +) AdamW launch of chunk i is issued on the current stream as soon as chunk i has arrived, i.e. it runs
    under the all-reduce of chunk i+1; only the last chunk's update is exposed.  `gcomm` (bf16 buffer, optional): the
    gradients travel in bf16 - cast, summed on the wire, cast back (DDP's bf16_compress_hook semantics)."""
    K = ops.get()
    g = fp.grad
    wire = g
    if gcomm is not None:
        K.cast(g[:fp.n_used], gcomm.dtype, out=gcomm)
        wire = gcomm
    works = [torch.distributed.all_reduce(wire[lo:hi], group=pg, async_op=True) for lo, hi, _ in plan]
    b1, b2 = betas
    sh_all = fp.shadow if shadow else None
    for w, (lo, hi, wd) in zip(works, plan):
        w.wait()                                        # the current stream waits for this chunk only
        if gcomm is not None:
            K.cast(gcomm[lo:hi], torch.float32, out=g[lo:hi])
        sh = sh_all[lo:hi] if sh_all is not None else None
        if clip > 0:
            K.grad_sumsq(g[lo:hi], blk2seg[lo // ALIGN:hi // ALIGN], n_seg, out=seg_sumsq)
            K.adamw_clip(fp.flat[lo:hi], g[lo:hi], m1[lo:hi], m2[lo:hi], sh, lr, b1, b2, eps, wd, step,
                         blk2seg[lo // ALIGN:], seg_sumsq, clip, grad_scale=gscale)
        else:
            K.adamw(fp.flat[lo:hi], g[lo:hi], m1[lo:hi], m2[lo:hi], sh, lr, b1, b2, eps, wd, step, grad_scale=gscale)


class ShardPlan:
    """Sharded optimizer state (ZeRO-1 style) over the flat buffers: rank r owns elements [r*S, (r+1)*S) of the used
    part, S a multiple of 64 (the buffers carry SHARD_PAD elements of slack for the rounding)."""

    def __init__(self, fp, world, rank, wd):
        S = -(-fp.n_used // (world * ALIGN)) * ALIGN
        self.S, self.padded, self.lo = S, S * world, rank * S
        if self.padded > fp.flat.numel():
            raise ValueError(f"flat buffer of {fp.flat.numel()} elements cannot hold {world} shards of {S}")
        hi = min(self.lo + S, fp.n_used)
        self.hi_used = max(hi, min(self.lo, fp.n_used))          # end of this rank's part of the used range
        self.segs = []                                           # (lo, hi, weight decay) AdamW launches of this rank
        if self.lo < hi:
            if self.lo < fp.n_no_decay:
                self.segs.append((self.lo, min(hi, fp.n_no_decay), 0.0))
            if hi > fp.n_no_decay:
                self.segs.append((max(self.lo, fp.n_no_decay), hi, wd))


def sharded_reduce_and_update(fp, plan, pg, m1, m2, lr, betas, eps, step, gscale, clip, blk2seg, n_seg, seg_sumsq,
                              gcomm=None, shadow=False):
    """DistributedDataParallel's all-reduce + a replicated optimizer.step() (train/main.py:413,111) restated as
    reduce-scatter -> AdamW on the owned shard -> all-gather of the UPDATED weights in the dtype the forward reads.
    Same arithmetic (every parameter is updated exactly once, from the same summed gradient), less wire and less HBM:
    the gradient crosses NVLink once (reduce-scatter = half an all-reduce), the weights come back as bf16 shadows (a
    quarter of an all-reduce) instead of every rank repeating the whole AdamW (0.22 ms at 46 M parameters) on the
    full fp32 state.  The fp32 master weights and the moments of a rank are current for its own shard only:
    gather_full() before reading them (checkpoints).  Per-parameter clipping: partial squared norms over the shard,
    one tiny all-reduce of the [n_parameters] vector."""
    K = ops.get()
    g, S, lo = fp.grad, plan.S, plan.lo
    wire = g
    if gcomm is not None:
        K.cast(g[:plan.padded], gcomm.dtype, out=gcomm)
        wire = gcomm
    torch.distributed.reduce_scatter_tensor(wire[lo:lo + S], wire[:plan.padded], group=pg)
    if gcomm is not None:
        K.cast(gcomm[lo:lo + S], torch.float32, out=g[lo:lo + S])
    if clip > 0:
        if plan.hi_used > lo:
            K.grad_sumsq(g[lo:plan.hi_used], blk2seg[lo // ALIGN:plan.hi_used // ALIGN], n_seg, out=seg_sumsq)
        else:
            seg_sumsq.zero_()
        torch.distributed.all_reduce(seg_sumsq, group=pg)
    b1, b2 = betas
    sh_all = fp.shadow if shadow else None
    for a, b, wd in plan.segs:
        sh = sh_all[a:b] if sh_all is not None else None
        if clip > 0:
            K.adamw_clip(fp.flat[a:b], g[a:b], m1[a:b], m2[a:b], sh, lr, b1, b2, eps, wd, step, blk2seg[a // ALIGN:],
                         seg_sumsq, clip, grad_scale=gscale)
        else:
            K.adamw(fp.flat[a:b], g[a:b], m1[a:b], m2[a:b], sh, lr, b1, b2, eps, wd, step, grad_scale=gscale)
    pub = sh_all if sh_all is not None else fp.flat
    torch.distributed.all_gather_into_tensor(pub[:plan.padded], pub[lo:lo + S], group=pg)


def gather_full(plan, pg, *buffers):
    """Completes sharded buffers (fp32 master weights, AdamW moments) on every rank, in place."""
    for t in buffers:
        torch.distributed.all_gather_into_tensor(t[:plan.padded], t[plan.lo:plan.lo + plan.S], group=pg)


class FlatAdamW:
    def __init__(self, model, lr=1e-4, wd=1e-5, betas=(0.9, 0.999), eps=1e-8, process_group=None, sum_grads=False,
                 shard=False):
        self.model, self.lr, self.wd, self.betas, self.eps = model, lr, wd, betas, eps
        self.pg = process_group
        self.shard, self._shard_plan = bool(shard), None
        self.world = torch.distributed.get_world_size(process_group) if process_group is not None else 1
        self.grad_scale = 1.0 if sum_grads else 1.0 / self.world
        self.step_count = 0
        self.attached = False

    def attach(self):
        m = self.model
        unused = [n for n, p in m.named_parameters() if p.grad is None and p.requires_grad]
        fp = m.flat_params()
        fp.set_unused(unused)
        fp = m.flat_params()
        m._shadow_stale = True
        g = fp.grad_buffer()
        for n, p in m.named_parameters():
            if n not in fp.index:
                continue
            off, numel, shape = fp.index[n]
            p.grad = g[off:off + numel].view(shape) if n not in fp.unused else None
        for mod in m.modules():
            mod._grad_sink = True
        self.m1 = torch.zeros_like(fp.flat)
        self.m2 = torch.zeros_like(fp.flat)
        if m.compute_dtype == torch.bfloat16:
            fp.refresh_shadow()
            m._shadow_stale = False
            m._shadow_managed = True          # kept fresh by the AdamW kernel from here on
        g.zero_()
        self.blk2seg, self.seg_names = build_blk2seg(fp)
        self.seg_sumsq = torch.zeros(len(self.seg_names), dtype=torch.float32, device=fp.flat.device)
        self.attached = True

    def zero_grad(self):
        self.model.flat_params().grad.zero_()

    def accumulate(self, apply_update):
        """Gradient accumulation (`backprop_freq`, train/main_egoexo4d_distributed.py:175-181; the steps zero the flat
        gradient before every backward): adds this micro-batch's gradient to an accumulator; at an update the sum is
        moved back into the gradient buffer.  Returns True when the caller should reduce / clip / apply now."""
        if apply_update and not getattr(self, "_acc_pending", False):
            return True
        g = self.model.flat_params().grad
        if getattr(self, "_gacc", None) is None:
            self._gacc = torch.zeros_like(g)
        self._gacc.add_(g)
        self._acc_pending = True
        if not apply_update:
            return False
        g.copy_(self._gacc)
        self._gacc.zero_()
        self._acc_pending = False
        return True

    def flat_grad(self):
        fp = self.model.flat_params()
        return fp.grad[:fp.n_used]

    def allreduce(self):
        if self.world > 1:
            torch.distributed.all_reduce(self.flat_grad(), group=self.pg)

    def reduce_and_step(self, clip_grad=0.0, chunks=1):
        """allreduce() + step() as one pipelined pass (world > 1): chunked all-reduce, AdamW of chunk i under the
        all-reduce of chunk i+1.  Single process: step()."""
        if self.world == 1:
            return self.step(clip_grad=clip_grad)
        fp = self.model.flat_params()
        if self.shard:
            if self._shard_plan is None:
                self._shard_plan = ShardPlan(fp, self.world, torch.distributed.get_rank(self.pg), self.wd)
            self.step_count += 1
            sharded_reduce_and_update(fp, self._shard_plan, self.pg, self.m1, self.m2, self.lr, self.betas, self.eps,
                                      self.step_count, self.grad_scale, clip_grad, self.blk2seg, len(self.seg_names),
                                      self.seg_sumsq, shadow=self.model.compute_dtype == torch.bfloat16)
            return
        if getattr(self, "_plan", None) is None:
            self._plan = plan_chunks(fp, chunks, self.wd)
        self.step_count += 1
        reduce_and_update(fp, self._plan, self.pg, self.m1, self.m2, self.lr, self.betas, self.eps, self.step_count,
                          self.grad_scale, clip_grad, self.blk2seg, len(self.seg_names), self.seg_sumsq,
                          shadow=self.model.compute_dtype == torch.bfloat16)

    def grad_norms(self):
        """Per-parameter norms of the (mean) gradient in `seg_names` order: one launch (exg_grad_sumsq), no host
        synchronisation."""
        K = ops.get()
        K.grad_sumsq(self.flat_grad(), self.blk2seg, len(self.seg_names), out=self.seg_sumsq)
        return self.seg_sumsq.sqrt() * self.grad_scale

    def clip_per_param(self, clip_grad):
        """utils/train_utils.py:3-13 on the flat buffer: p.grad *= clip / (||p.grad|| + 1e-6) where that is < 1, per
        parameter, in place (two launches instead of ~170 norm / .item() / mul_ triples).  Returns the norms."""
        norms = self.grad_norms()
        ops.get().grad_clip_scale(self.flat_grad(), self.blk2seg, self.seg_sumsq, clip_grad, self.grad_scale)
        return norms

    def step(self, clip_grad=0.0):
        """Two AdamW launches (no-decay / decay segment).  clip_grad > 0: the reference's per-parameter clipping is
        folded into the optimizer's read of the gradient (exg_grad_sumsq + exg_adamw_clip; .grad itself is left
        unscaled)."""
        fp = self.model.flat_params()
        K = ops.get()
        self.step_count += 1
        b1, b2 = self.betas
        shadow = fp.shadow if self.model.compute_dtype == torch.bfloat16 else None
        if clip_grad > 0:
            K.grad_sumsq(self.flat_grad(), self.blk2seg, len(self.seg_names), out=self.seg_sumsq)
        for lo, hi, wd in ((0, fp.n_no_decay, 0.0), (fp.n_no_decay, fp.n_used, self.wd)):
            if hi > lo:
                sh = shadow[lo:hi] if shadow is not None else None
                if clip_grad > 0:
                    K.adamw_clip(fp.flat[lo:hi], fp.grad[lo:hi], self.m1[lo:hi], self.m2[lo:hi], sh, self.lr, b1, b2,
                                 self.eps, wd, self.step_count, self.blk2seg[lo // ALIGN:], self.seg_sumsq, clip_grad,
                                 grad_scale=self.grad_scale)
                else:
                    K.adamw(fp.flat[lo:hi], fp.grad[lo:hi], self.m1[lo:hi], self.m2[lo:hi], sh, self.lr, b1, b2,
                            self.eps, wd, self.step_count, grad_scale=self.grad_scale)

    def sync_full(self):
        """shard=True: completes the fp32 master weights and the moments on every rank (collective)."""
        if self._shard_plan is not None:
            gather_full(self._shard_plan, self.pg, self.model.flat_params().flat, self.m1, self.m2)

    def state_dict(self):
        """torch.optim.AdamW layout over the reference's two parameter groups (see optim_state_dict)."""
        self.sync_full()
        return optim_state_dict(self.model, self.m1, self.m2, self.step_count, self.lr, self.wd, self.betas, self.eps)

    def load_state_dict(self, sd):
        self.step_count, self.lr, self.wd = load_optim_state_dict(self.model, self.m1, self.m2, sd)
