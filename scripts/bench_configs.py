"""Secondary BASELINE.json configurations (the headline configs[1] is bench.py):

  --exo    configs[2]: ExoGroundingTransformer keystep grounding, Ego-Exo4D-shaped synthetic features
           (T=1024 video tokens, 768-d video, 4096-d narrations), forward + grounding/InfoNCE loss + backward.
  --sweep  configs[4]: inference alignment-logit sweep, TAN E6D6 joint + dual similarity forward
           (get_text_visual_sim_joint / _dual, eval/eval_zeroshot_align.py's per-window call) on HTM-Align-shaped
           inputs, T = 64 -> 4096 seconds per window, 32 sentences, 80 videos in total per T.
  --align  sliding-window alignment (exoground_b200.eval.align): 80 synthetic videos of ~15 min, seq_len 64,
           batched windows vs the reference's window loop.

Device timings with CUDA events after warm-up; one JSON line per measurement.
"""
import argparse
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from exoground_b200 import ops                                    # noqa: E402
from exoground_b200.ops import lib                                # noqa: E402
from exoground_b200.utils import synth                            # noqa: E402


def timed(fn, warmup, steps):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def breakdown(fn):
    fn(); torch.cuda.synchronize()
    lib.profile_begin()
    fn()
    prof = lib.profile_end()
    tot = sum(v["ms"] for v in prof.values())
    out = {}
    for k, v in prof.items():
        tensor = k.startswith("gemm") or k.startswith("attn")
        rate = v["work"] / (v["ms"] * 1e-3) / (1e12 if tensor else 1e9) if v["ms"] > 0 else 0.0
        out[k] = {"ms": round(v["ms"], 3), "share": round(v["ms"] / tot, 3), ("tflops" if tensor else "gbs"): round(rate, 1)}
    return out


def run_exo(a):
    import os
    import torch.distributed as dist
    from exoground_b200.model import ExoGroundingTransformer
    from exoground_b200.train.step_exo import ExoTrainStep
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", 1), ("RANK", 0), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    pg = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
        pg = dist.group.WORLD
    B, T, N, L = a.batch or 16, 1024, 64, a.layers             # SURVEY cfg-3: E6D6, T=1024, N=64, 768-d / 4096-d
    torch.manual_seed(0); np.random.seed(0)
    m = ExoGroundingTransformer(L, L, video_embed_dim=768, text_embed_dim=4096, use_distill_nce_loss=a.distill,
                                random_pos_start=0, compute_dtype=torch.bfloat16)   # T = the whole 1024-row table
    b = synth.exo_batch(B=B, T=T, N=N, d_video=768, d_text=4096, d_ego=768 if a.distill else None, seed=888 + rank)
    b = {k: (v.pin_memory() if torch.is_tensor(v) else v) for k, v in b.items()}
    args = synth.exo_args(use_distill_nce_loss=a.distill, same_view_negative=a.distill)
    st = ExoTrainStep(m, args, device=f"cuda:{local}", process_group=pg, use_graph=not a.no_graph,
                      shard_optimizer=not a.replicated_optimizer)

    def step():
        st.step(b, next_batch=b)      # double-buffered input: every step contains one full H2D copy (copy stream)

    ms = timed(step, a.warmup, a.steps)
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    graph, st.graph = st.graph, None           # per-class breakdown from one eager step
    ug, st.use_graph = st.use_graph, False
    bd = breakdown(step)
    st.graph, st.use_graph = graph, ug
    # algorithmic train FLOPs per clip (SURVEY §8d formulas): unimodal video + text encoders, joint encoder, decoder
    C, F = 512, 2048
    per_tok = lambda S: 2 * C * 3 * C + 2 * C * C + 4 * C * F + 4 * S * C
    dec_tok = per_tok(N) + (2 * C * C + 2 * C * C + 4 * T * C)          # self block + cross q / out proj + cross attn
    fwd = (T * 2 * 768 * C + N * 2 * 4096 * C + L * T * per_tok(T) + L * N * per_tok(N) + L * (T + N) * per_tok(T + N)
           + L * (N * dec_tok + T * 2 * C * 2 * C) + (T * 2 * C * 768 if a.distill else 0))
    gf = 3 * fwd / 1e9
    if rank == 0:
        cps = B * world / (ms * 1e-3)
        print(json.dumps({"config": f"ExoGround E{L}D{L} keystep grounding, full train step: H2D + fwd + grounding"
                                    f"{' / view-invariant InfoNCE' if a.distill else ''} loss + bwd + grad all-reduce + "
                                    f"AdamW (BASELINE configs[2])",
                          "T": T, "N": N, "n_gpus": world, "per_gpu_batch": B, "dtype": "bf16",
                          "ms_per_step": round(ms, 3), "clips_per_s": round(cps, 1),
                          "train_gflop_per_clip": round(gf, 1), "model_tflops_per_gpu": round(cps / world * gf / 1e3, 1),
                          "h2d_bytes_per_step": st.h2d_bytes, "cuda_graph": bool(ug),
                          "breakdown": bd}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_sweep(a):
    from exoground_b200.model import TemporalAligner
    torch.manual_seed(0)
    m = TemporalAligner(6, 6, use_alignability_head=1, video_embed_dim=1024, text_embed_dim=4096, random_pos_start=0,
                        compute_dtype=torch.bfloat16).cuda().eval()
    N, total = 32, 80
    for T in (64, 128, 256, 512, 1024, 2048, 4096):
        B = max(1, min(total, 8192 // T))            # videos per forward; 80 videos = ceil(80 / B) forwards
        video = torch.randn(B, T, 1024, device="cuda")
        text = torch.randn(B, N, 4096, device="cuda")

        def fwd():
            with torch.no_grad():
                m.get_text_visual_sim_joint(video, text)
                m.get_text_visual_sim_dual(video, text)

        ms = timed(fwd, a.warmup, a.steps)
        bd = breakdown(fwd)
        from exoground_b200.eval.graphed import GraphedSimilarity
        gs = GraphedSimilarity(m, B, T, N, 1024, 4096)
        with torch.no_grad():
            want_j = m.get_text_visual_sim_joint(video, text)
        got_j, _ = gs(video, text)
        assert torch.equal(got_j, want_j), "graph replay differs from the eager helpers"
        ms_g = timed(lambda: gs(video, text), a.warmup, a.steps)
        print(json.dumps({"config": "TAN E6D6 alignment-logit sweep (BASELINE configs[4])", "T": T, "N": N,
                          "videos_per_forward": B, "ms_per_forward": round(ms, 3),
                          "ms_per_forward_graph": round(ms_g, 3), "videos_per_s_graph": round(B / (ms_g * 1e-3), 1),
                          "videos_per_s": round(B / (ms * 1e-3), 1),
                          "ms_80_videos": round(ms * int(np.ceil(total / B)), 2),
                          "attn_fwd": bd.get("attn_fwd"), "gemm_tc": bd.get("gemm_tc")}), flush=True)


def run_align(a):
    from exoground_b200.eval import align
    from exoground_b200.model import TemporalAligner
    torch.manual_seed(0)
    m = TemporalAligner(6, 6, use_alignability_head=0, video_embed_dim=1024, text_embed_dim=4096, random_pos_start=0,
                        compute_dtype=torch.bfloat16).cuda().eval()
    items = []
    for i in range(a.videos):
        it = synth.htm_align_item(vlen=600 + 40 * (i % 16), K=80 + 5 * (i % 9), d_video=1024, d_text=4096, seed=100 + i)
        it["video"], it["text"] = it["video"].cuda(), it["text"].cuda()
        items.append(it)
    for batched in (True, False):
        align.test_alignment(m, items[:2], 64, False, batched)          # warm-up
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = align.test_alignment(m, items, 64, False, batched)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        nwin = sum(len(align.window_plan(it["video"].shape[1], 64, it["start"], it["end"], it["aligned"])) for it in items)
        print(json.dumps({"config": "sliding-window alignment, seq_len 64 (eval_zeroshot_align.py:97-252)",
                          "videos": len(items), "windows": nwin, "mode": "batched" if batched else "window loop",
                          "seconds": round(dt, 3), "videos_per_s": round(len(items) / dt, 1),
                          "Recall": res["Recall"], "AUC": res["AUC"]}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--exo", action="store_true")
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--align", action="store_true")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--layers", type=int, default=6, help="--exo: encoder = decoder depth (SURVEY cfg-3: 6)")
    ap.add_argument("--distill", action="store_true", help="--exo: with the ego-distillation InfoNCE (768-d ego features)")
    ap.add_argument("--no-graph", action="store_true", help="--exo: eager step instead of the CUDA-graph replay")
    ap.add_argument("--replicated-optimizer", action="store_true",
                    help="--exo, N>1: all-reduce + replicated AdamW instead of the sharded optimizer")
    ap.add_argument("--videos", type=int, default=80)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    assert torch.cuda.is_available() and ops.get().name == "cuda"
    if a.exo:
        run_exo(a)
    if a.sweep:
        run_sweep(a)
    if a.align:
        run_align(a)


if __name__ == "__main__":
    main()
