"""The TAN training loop of the reference (train/main.py:36-157), repaired and runnable on the drop-in.

The loop as shipped cannot run: it feeds `input_data['video']` / `['padding_mask']` as BOTH the video and the text
tensors (:53-56; the collate dict of data/loader_egoexo4d_tan.py:333-341 carries the sentences under
'narration_features' / 'narration_padding_mask'), and it calls `evaluate_downstream` (HTM-Align + YouCook2 files) after
every iteration (:151).  This is the same loop with those two lines repaired - sentences read from their own keys, the
downstream evaluation behind `args.eval_every` - and otherwise the reference's order of operations: autocast forward
(:75-81), the EMA-target forward for 'cotrain' (:82-89), get_loss (:91-98), scaled backward (:106), every
`backprop_freq` iterations unscale -> per-parameter clip -> optimizer step -> scaler update -> zero_grad (:107-113)
and the momentum update (:116-117), logging every 5 iterations (:120-126), lr_scheduler.step(iteration) (:135), the
runtime checkpoint with `'optimizer': optimizer.state_dict()` (:139-150).

Two ways to run it:
  * generic (`fused_step=None`): torch autograd over the drop-in model + any torch optimizer - a drop-in for the
    reference loop, e.g. under DistributedDataParallel;
  * fused (`fused_step=TanTrainStep(...)`): forward + loss + backward replayed as one CUDA graph, gradient all-reduce,
    fused clip + AdamW (+ EMA) on the flat buffers; `optimizer` / `grad_scaler` are then unused (bf16 needs no loss
    scaling) and the checkpoint's optimizer state comes from `fused_step.state_dict()` in torch.optim.AdamW's layout.
"""
import os
import time

import torch

from ..ops import get as _ops
from . import loss as exg_loss


def batch_from_collate(input_data):
    """Collate dict of data/loader_egoexo4d_tan.py:333-341 -> the tensors the step consumes (the repair of
    train/main.py:53-56)."""
    return {"video": input_data["video"], "video_padding_mask": input_data["padding_mask"].bool(),
            "text": input_data["narration_features"], "text_padding_mask": input_data["narration_padding_mask"].bool(),
            "start": input_data["start"], "end": input_data["end"],
            "abs_text_start": input_data.get("abs_text_start"), "abs_text_end": input_data.get("abs_text_end")}


def clip_gradients(model, clip_grad):
    """utils/train_utils.py:3-13 without its one `.item()` per parameter (same arithmetic, no host sync)."""
    ps = [p for p in model.parameters() if p.grad is not None]
    if not ps:
        return []
    norms = torch._foreach_norm([p.grad for p in ps])
    coef = [(clip_grad / (n + 1e-6)).clamp(max=1.0) for n in norms]
    torch._foreach_mul_([p.grad for p in ps], coef)
    return norms


def train(loader, model, optimizer, lr_scheduler, grad_scaler, device, epoch, args, fused_step=None):
    """One epoch; returns the average loss (train/main.py:36-157).  `args`: model, backprop_freq, clip_grad, iteration,
    rank, runtime_save_iter, model_path, train_plotter (optional), eval_every / evaluate_fn (optional) + the fields
    get_loss reads."""
    rank = getattr(args, "rank", 0)
    plotter = getattr(args, "train_plotter", None)
    total, count = 0.0, 0
    model.train()
    tic = time.time()
    if fused_step is None:
        optimizer.zero_grad()
    elif lr_scheduler is not None:
        fused_step.lr = lr_scheduler.get_last_lr()[0]
    core = model.module if hasattr(model, "module") else model
    on_cuda = torch.device(device).type == "cuda" and _ops().name == "cuda"
    for idx, input_data in enumerate(loader):
        end = time.time()
        b = batch_from_collate(input_data)
        B = b["video"].shape[0]
        if fused_step is not None:
            # train/main.py:107: the optimizer steps on iterations 0, k, 2k, ... of the epoch; in between the gradients add up
            loss_val = fused_step.step(b, apply_update=(idx % args.backprop_freq == 0))
            loss_dict = {"loss": torch.tensor(loss_val)}
        else:
            video = b["video"].to(device, non_blocking=True)
            text = b["text"].to(device, non_blocking=True)
            vmask = b["video_padding_mask"].to(device, non_blocking=True)
            tmask = b["text_padding_mask"].to(device, non_blocking=True)
            abs_text_pos = None
            if b["abs_text_start"] is not None:
                abs_text_pos = exg_loss.get_text_pos(b["abs_text_start"], b["abs_text_end"], device=device)
            with torch.autocast("cuda", enabled=on_cuda):
                logits = model(video, text, video_padding_mask=vmask, lang_padding_mask=tmask, text_timestamp=None,
                               abs_text_pos=abs_text_pos)
                if args.model in ("cotrain",):
                    with torch.no_grad():
                        ema = core.forward_from_ema(video, text, video_padding_mask=vmask, lang_padding_mask=tmask,
                                                    text_timestamp=None, abs_text_pos=abs_text_pos)
                    logits = {**logits, **{f"ema-{k}": v for k, v in ema.items()}}
                loss_dict = exg_loss.get_loss(input_data={"start": b["start"], "end": b["end"]}, video_seq=video,
                                              text_embed=text, video_padding_mask=vmask, text_padding_mask=tmask,
                                              logits=logits, args=args, abs_text_pos=abs_text_pos)
            loss = loss_dict["loss"]
            grad_scaler.scale(loss).backward()
            if idx % args.backprop_freq == 0:
                grad_scaler.unscale_(optimizer)
                if args.clip_grad > 0:
                    clip_gradients(model, args.clip_grad)
                grad_scaler.step(optimizer)
                grad_scaler.update()
                optimizer.zero_grad()
                if args.model in ("cotrain",):
                    core._momentum_update()
            loss_val = float(loss.detach())
        if loss_val == loss_val and abs(loss_val) != float("inf"):
            total += loss_val * B
            count += B
        if rank == 0 and plotter is not None and args.iteration % 5 == 0:
            for k, v in loss_dict.items():
                plotter.add_data(f"train/{k}", float(v), args.iteration)
            if lr_scheduler is not None:
                plotter.add_data("train/lr", lr_scheduler.get_last_lr()[0], args.iteration)
            plotter.add_data("device/sps", 1 / max(time.time() - end, 1e-9), args.iteration)
        if lr_scheduler is not None:
            lr_scheduler.step(args.iteration)
            if fused_step is not None:
                fused_step.lr = lr_scheduler.get_last_lr()[0]
        args.iteration += 1
        if rank == 0 and getattr(args, "runtime_save_iter", 0) and args.iteration % args.runtime_save_iter == 0:
            opt_state = fused_step.state_dict() if fused_step is not None else optimizer.state_dict()
            torch.save({"epoch": epoch, "state_dict": model.state_dict(), "best_acc": 1e5, "optimizer": opt_state,
                        "iteration": args.iteration}, os.path.join(args.model_path, "runtime.pth.tar"))
        every = getattr(args, "eval_every", 0)
        if every and args.iteration % every == 0 and getattr(args, "evaluate_fn", None) is not None:
            for k, v in args.evaluate_fn(model, device, args).items():      # train/main.py:151-153, on a schedule
                if plotter is not None:
                    plotter.add_data(f"local/{k}", float(v), args.iteration)
            model.train()
    if rank == 0 and plotter is not None:
        plotter.add_data("global/loss", total / max(count, 1), epoch)
    args.epoch_seconds = time.time() - tic
    return total / max(count, 1)
