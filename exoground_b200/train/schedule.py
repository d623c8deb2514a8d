"""Learning-rate schedule of the reference trainers: linear warm-up over `warmup_iterations` (1000) iterations, then half
a cosine down to zero at `epochs * iter_per_epoch` (train/main_egoexo4d_distributed.py:832-843, train/main.py:487-499 -
`lr_schedule_fn` wrapped in `torch.optim.lr_scheduler.LambdaLR`, stepped with the GLOBAL iteration: `:201`,
`lr_scheduler.step(args.iteration)`).

The fused steps (TanTrainStep / ExoTrainStep) take `lr` as a plain attribute and own no torch optimizer, so
`WarmupCosine` is the same schedule behind the three calls the loops make on a LambdaLR: `step(iteration)`,
`get_last_lr()`, `state_dict()` / `load_state_dict()` - no dummy optimizer needed."""
import math


def lr_multiplier(iteration, iter_per_epoch, epochs, warmup_iterations=1000):
    """`lr_schedule_fn` (main_egoexo4d_distributed.py:833-839): the factor applied to the base lr at `iteration`."""
    if iteration < warmup_iterations:
        return iteration / warmup_iterations
    return 0.5 * (1.0 + math.cos(math.pi * (iteration - warmup_iterations)
                                 / (epochs * iter_per_epoch - warmup_iterations)))


class WarmupCosine:
    def __init__(self, base_lr, iter_per_epoch, epochs, warmup_iterations=1000, targets=(), iteration=0):
        """`targets`: objects with an `lr` attribute (TanTrainStep, ExoTrainStep.opt / FlatAdamW) updated on every step()."""
        self.base_lr, self.iter_per_epoch, self.epochs = float(base_lr), int(iter_per_epoch), int(epochs)
        self.warmup_iterations = int(warmup_iterations)
        self.targets = list(targets)
        self.last_epoch = -1
        self.step(iteration)                       # LambdaLR evaluates the schedule once at construction

    def _lr(self, it):
        return self.base_lr * lr_multiplier(it, self.iter_per_epoch, self.epochs, self.warmup_iterations)

    def step(self, iteration=None):
        """LambdaLR.step(epoch=iteration): closed form at that iteration (the reference always passes it); without an
        argument the counter advances by one."""
        self.last_epoch = self.last_epoch + 1 if iteration is None else int(iteration)
        lr = self._lr(self.last_epoch)
        for t in self.targets:
            t.lr = lr
        return lr

    def get_last_lr(self):
        return [self._lr(self.last_epoch)]

    def state_dict(self):
        return {"last_epoch": self.last_epoch, "base_lrs": [self.base_lr]}

    def load_state_dict(self, sd):
        self.step(sd["last_epoch"])
