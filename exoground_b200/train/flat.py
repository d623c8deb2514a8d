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
        self.attached = True

    def zero_grad(self):
        self.model.flat_params().grad.zero_()

    def flat_grad(self):
        fp = self.model.flat_params()
        return fp.grad[:fp.n_used]

    def allreduce(self):
        if self.world > 1:
            torch.distributed.all_reduce(self.flat_grad(), group=self.pg)

    def clip_per_param(self, clip_grad):
        """utils/train_utils.py:3-13 on the flat buffer: p.grad *= min(1, clip / (||p.grad|| + 1e-6)) per parameter;
        returns the norms as one device tensor (no host synchronisation)."""
        views = [p.grad for _, p in self.model.named_parameters() if p.grad is not None]
        norms = torch.stack(torch._foreach_norm(views)) * self.grad_scale      # of the MEAN gradient, as under DDP
        coef = (clip_grad / (norms + 1e-6)).clamp(max=1.0)
        torch._foreach_mul_(views, list(coef.unbind()))
        return norms

    def step(self):
        fp = self.model.flat_params()
        K = ops.get()
        self.step_count += 1
        b1, b2 = self.betas
        shadow = fp.shadow if self.model.compute_dtype == torch.bfloat16 else None
        for lo, hi, wd in ((0, fp.n_no_decay, 0.0), (fp.n_no_decay, fp.n_used, self.wd)):
            if hi > lo:
                K.adamw(fp.flat[lo:hi], fp.grad[lo:hi], self.m1[lo:hi], self.m2[lo:hi],
                        shadow[lo:hi] if shadow is not None else None, self.lr, b1, b2, self.eps, wd,
                        self.step_count, grad_scale=self.grad_scale)
