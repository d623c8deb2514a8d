"""Parameter storage for the B200 path: one flat fp32 buffer per model (+ a flat bf16 shadow copy
refreshed by one cast kernel, + a flat fp32 gradient buffer for the fused optimizer / all-reduce).

`nn.Parameter`s stay ordinary parameters (same names as the reference, so reference checkpoints
load and torch DDP / optimizers still work); only their `.data` is re-pointed at views of the flat
buffer.  Offsets are padded to 64 elements so every view is 128-byte aligned (TMA needs 16 B).
"""
import os

import torch

from .. import ops

ALIGN = 64


def default_compute_dtype():
    v = os.environ.get("EXG_COMPUTE_DTYPE", "bf16").lower()
    return torch.float32 if v in ("fp32", "float32", "f32") else torch.bfloat16


SHARD_PAD = 64 * 64          # slack behind the flat buffers: up to 64 ranks x one 64-element alignment unit


def is_no_decay(name):
    """The reference's optimiser policy (train/main.py:352): no weight decay for LayerNorms and biases."""
    name = "." + name
    return any(k in name for k in (".ln_", ".bias", ".logit_scale", ".entropy_scale"))


def own_named_parameters(module):
    """named_parameters() minus those that live under a nested ExgModule (e.g. the frozen `vi_encoder` inside
    keystep_ground_model.GroundingModel), which keeps its own flat buffer."""
    nested = [n + "." for n, m in module.named_modules() if n and isinstance(m, ExgModule)]
    return [(n, p) for n, p in module.named_parameters() if not any(n.startswith(pre) for pre in nested)]


class FlatParams:
    def __init__(self, module):
        self.module = module
        self.flat = None
        self.shadow = None
        self.grad = None
        self.index = {}          # name -> (offset, numel, shape)
        self.n_no_decay = 0      # elements [0, n_no_decay) take no weight decay
        self.n_used = 0          # elements [n_no_decay, n_used) take weight decay; [n_used, end) never get a gradient
        self.unused = set()      # parameter names that receive no gradient (torch.optim skips them: grad is None)
        self.layer_ranges = {}   # "<stack>.resblocks.<i>." -> (lo, hi): that block's weight-decayed matrices, contiguous
        self.n_bucketed = 0      # elements [n_bucketed, n_used) are covered by layer_ranges (overlapped all-reduce)
        self._ptr_check = []

    def set_unused(self, names):
        """Move parameters that never receive a gradient (reference: `mlp`, `text_temporal_pos_embed`, ... —
        the reason for DDP's find_unused_parameters=True, train/main.py:413) to the tail of the flat buffer,
        which the fused optimizer and the gradient all-reduce skip (torch.optim skips grad=None parameters)."""
        names = set(names)
        if names != self.unused:
            self.unused = names
            self.flat = None

    def _layout(self):
        def cls(n):
            return 2 if n in self.unused else (0 if is_no_decay(n) else 1)

        def order(n):
            # decayed matrices of transformer blocks go last inside the decay segment, block by block, so a
            # trainer can all-reduce one block's gradients as one contiguous range while backward continues
            return (cls(n), 1 if (cls(n) == 1 and ".resblocks." in "." + n) else 0)
        named = own_named_parameters(self.module)
        named.sort(key=lambda np_: order(np_[0]))
        off = 0
        index = {}
        ends = [0, 0, 0]
        ranges = {}
        for n, p in named:
            index[n] = (off, p.numel(), tuple(p.shape))
            lo = off
            off += (p.numel() + ALIGN - 1) // ALIGN * ALIGN
            ends[cls(n)] = off
            if order(n) == (1, 1):
                pre = n[:n.index(".", n.index("resblocks.") + len("resblocks.")) + 1]
                r = ranges.get(pre)
                ranges[pre] = (lo, off) if r is None else (r[0], off)
        n_no_decay = ends[0]
        n_used = max(ends[0], ends[1])
        self._ranges = ranges
        self._n_bucketed = min((r[0] for r in ranges.values()), default=n_used)
        return named, index, off, n_no_decay, n_used

    def ensure(self):
        """(Re)build the flat buffer if parameters were moved / replaced since the last call."""
        if self.flat is not None and all(p.data_ptr() == ptr for p, ptr in self._ptr_check):
            return False
        named, index, total, n_no_decay, n_used = self._layout()
        dev = named[0][1].device
        # + SHARD_PAD never-used elements: a sharded optimizer (train/flat.py::ShardPlan) rounds the used part up to a
        # whole number of equal, 64-aligned shards and reduce-scatters / all-gathers that padded range in place
        flat = torch.zeros(total + SHARD_PAD, dtype=torch.float32, device=dev)
        self._ptr_check = []
        for n, p in named:
            off, numel, shape = index[n]
            view = flat[off:off + numel].view(shape)
            view.copy_(p.data)
            p.data = view
            self._ptr_check.append((p, p.data_ptr()))
        self.flat, self.index, self.n_no_decay, self.n_used = flat, index, n_no_decay, n_used
        self.layer_ranges, self.n_bucketed = self._ranges, self._n_bucketed
        self.shadow = None
        self.grad = None
        return True

    def refresh_shadow(self):
        """bf16 copy of every weight: one cast kernel over the flat buffer."""
        self.ensure()
        if self.shadow is None:
            self.shadow = torch.empty(self.flat.numel(), dtype=torch.bfloat16, device=self.flat.device)
        ops.get().cast(self.flat, torch.bfloat16, out=self.shadow)
        return self.shadow

    def shadow_view(self, name):
        off, numel, shape = self.index[name]
        return self.shadow[off:off + numel].view(shape)

    def grad_buffer(self):
        self.ensure()
        if self.grad is None:
            self.grad = torch.zeros_like(self.flat)
        return self.grad

    def grad_views(self):
        g = self.grad_buffer()
        return {n: g[off:off + numel].view(shape) for n, (off, numel, shape) in self.index.items()}


class ExgModule(torch.nn.Module):
    """Base of the drop-in model classes: owns the FlatParams and the compute dtype."""

    def _exg_init(self, compute_dtype=None):
        object.__setattr__(self, "_flat", FlatParams(self))
        self.compute_dtype = compute_dtype or default_compute_dtype()
        self._shadow_stale = True

    def set_compute_dtype(self, dtype):
        if dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("compute dtype must be torch.float32 or torch.bfloat16")
        self.compute_dtype = dtype
        return self

    def flat_params(self):
        self._flat.ensure()
        return self._flat

    def _weights(self):
        """Returns wT(name) -> weight tensor in compute dtype.  Refreshes the bf16 shadow when training
        (weights change every step) or when it has never been built."""
        T = self.compute_dtype
        fp = self._flat
        rebuilt = fp.ensure()
        if T == torch.float32:
            params = dict(self.named_parameters())
            return lambda name: params[name]
        managed = getattr(self, "_shadow_managed", False)     # a trainer's fused AdamW keeps the shadow current
        refresh = rebuilt or fp.shadow is None or self._shadow_stale or \
            (not managed and (self.training or torch.is_grad_enabled()))
        if not refresh and not managed:
            # evaluation right after an external optimizer step (the reference's evaluate() follows train() with
            # torch.optim.AdamW): in-place updates bump the parameters' version counters
            refresh = self._param_version() != getattr(self, "_shadow_version", None)
        if refresh:
            fp.refresh_shadow()
            self._shadow_stale = False
            self._shadow_version = self._param_version()
        return fp.shadow_view

    def _param_version(self):
        return sum(p._version for p in self.parameters())

    def load_state_dict(self, state_dict, strict=True, **kw):
        # accept checkpoints saved from a DDP wrapper (keys carry "module.":
        # train/main_egoexo4d_distributed.py:716-717,870)
        if state_dict and all(k.startswith("module.") for k in state_dict):
            state_dict = {k[len("module."):]: v for k, v in state_dict.items()}
        self._shadow_stale = True
        res = super().load_state_dict(state_dict, strict=strict, **kw)
        # a trainer-managed shadow is only refreshed by the AdamW kernel, and a CUDA-graph replay never goes through
        # _weights(): refresh it now, or the next step would differentiate the OLD bf16 weights
        if self._flat.flat is not None and self._flat.shadow is not None:
            self._flat.ensure()
            self._flat.refresh_shadow()
            self._shadow_stale = False
        return res
