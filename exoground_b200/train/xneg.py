"""Cross-GPU negatives ("Mode B"; BASELINE.json north star): the all-pairs text-by-time logits of the TAN
loss (reference train/loss.py:234-278 over `einsum("astc,bkc->astbk")`, tan_model.py:112-113) taken over the GLOBAL
batch while every rank keeps only its own clips' rows.

Per rank r of W (B clips each, Bg = W * B):
  * the L2-normalised text features (dual: [B*N, C]; joint: [D, B*N, C]) are all-gathered (rank-major, so the
    global column clip of local clip a is a + r * B) -> the rank computes the row block [S, B*T, Bg*N] of the
    global logits with the same similarity GEMM; the score rows of a clip never leave its rank;
  * `exg_nce_fwd(..., Bc=Bg, clip0=r*B)` gives complete row exp-sums and PARTIAL column exp-sums; one all-reduce of
    the [4, S, Bg*N]-sized column sums completes them (logits are cosines, so the fixed shift of the kernels makes
    the partial sums directly addable: no running max to reconcile);
  * loss = sum over ranks of (own rows' terms + own columns' terms); every rank differentiates the GLOBAL loss
    with respect to its row block (`exg_nce_bwd` with the global column sums), the similarity GEMM's backward
    yields gradients for ALL text columns, and the all-gather's backward reduce-scatters them to their owners;
  * parameter gradients are then SUMMED over ranks (global mean loss => sum of per-rank partial gradients, not
    DDP's mean).

`Group` wraps the few collectives this needs so the same loss code runs with world size 1 (all identities), on NCCL
(also under CUDA-graph capture) and on gloo (the CPU tests).
"""
import torch
import torch.distributed as dist


class Group:
    def __init__(self, process_group=None):
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if process_group is not None else 1
        self.rank = dist.get_rank(process_group) if process_group is not None else 0

    def gather0(self, x):
        """Concatenate x over ranks along dim 0 (rank-major)."""
        if self.world == 1:
            return x
        x = x.contiguous()
        as_bool = x.dtype == torch.bool
        if as_bool:
            x = x.to(torch.uint8)
        out = torch.empty((self.world * x.shape[0],) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, x, group=self.pg)
        return out.bool() if as_bool else out

    def sum_(self, x):
        """In-place sum over ranks."""
        if self.world > 1:
            dist.all_reduce(x, group=self.pg)
        return x

    def local0(self, x, n):
        """This rank's n-row slice of a rank-major dim-0 concatenation."""
        return x[self.rank * n:(self.rank + 1) * n]


class AllGatherRowsFn(torch.autograd.Function):
    """x [rows, C] or [S, rows, C] -> [W*rows, C] / [S, W*rows, C] (rank-major along the row axis).
    Backward: reduce-scatter (sum) of the gradient, i.e. each rank receives the sum over ranks of the gradient of
    ITS rows."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group, ctx.three = group, x.dim() == 3
        W = group.world
        x = x.contiguous()
        buf = torch.empty((W,) + tuple(x.shape), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(buf.view(W * x.shape[0], *x.shape[1:]), x, group=group.pg)
        if ctx.three:                                     # [W, S, rows, C] -> [S, W*rows, C]
            S, rows, C = x.shape
            return buf.permute(1, 0, 2, 3).reshape(S, W * rows, C)
        return buf.view(W * x.shape[0], x.shape[1])

    @staticmethod
    def backward(ctx, g):
        group = ctx.group
        W = group.world
        if ctx.three:
            S, rows_g, C = g.shape
            rows = rows_g // W
            g = g.reshape(S, W, rows, C).permute(1, 0, 2, 3).contiguous()      # [W, S, rows, C]
            out = torch.empty((S, rows, C), dtype=g.dtype, device=g.device)
        else:
            rows_g, C = g.shape
            rows = rows_g // W
            g = g.contiguous()
            out = torch.empty((rows, C), dtype=g.dtype, device=g.device)
        if g.device.type == "cuda":
            dist.reduce_scatter_tensor(out, g.view(W * out.shape[0], *out.shape[1:]), group=group.pg)
        else:                                             # gloo has no reduce_scatter: all-reduce + slice
            dist.all_reduce(g, group=group.pg)
            out.copy_(g.view(W, *out.shape)[group.rank])
        return out, None


def gather_rows(x, group):
    return x if group is None or group.world == 1 else AllGatherRowsFn.apply(x, group)
