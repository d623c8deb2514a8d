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


class FlatAdamW:
    def __init__(self, model, lr=1e-4, wd=1e-5, betas=(0.9, 0.999), eps=1e-8, process_group=None, sum_grads=False):
        self.model, self.lr, self.wd, self.betas, self.eps = model, lr, wd, betas, eps
        self.pg = process_group
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

    def flat_grad(self):
        fp = self.model.flat_params()
        return fp.grad[:fp.n_used]

    def allreduce(self):
        if self.world > 1:
            torch.distributed.all_reduce(self.flat_grad(), group=self.pg)

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

    def state_dict(self):
        """torch.optim.AdamW layout over the reference's two parameter groups (see optim_state_dict)."""
        return optim_state_dict(self.model, self.m1, self.m2, self.step_count, self.lr, self.wd, self.betas, self.eps)

    def load_state_dict(self, sd):
        self.step_count, self.lr, self.wd = load_optim_state_dict(self.model, self.m1, self.m2, sd)
