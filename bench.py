#!/usr/bin/env python
"""bench.py — training clips/sec of the TAN E6D6 alignment hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Own arm: one process per GPU (torchrun for N>1), per-GPU batch fixed (weak scaling).  A step is one full
training step of TemporalAligner E6D6 (T=64, N=32 sentences, 4096-d features, bf16 compute, fp32 master
weights): forward -> MIL-NCE + alignability loss -> backward -> gradient all-reduce -> fused AdamW.
  value : device-resident inputs, K steps between barrier+synchronize, CUDA events, max over ranks;
  e2e   : the same K steps through TanTrainStep.step(batch) with pinned-host inputs (H2D copies, host-built
          target masks and the loss read-back inside the timed region).
  roofline / breakdown : per-kernel-class CUDA-event timings of one profiled step (exg_profile_*).
  sustained : the same device-resident step repeated for >= 3 s with its own clock record (burst vs sustained).
  cpu_baseline : the reference's own modules (oracle/_ref archive; kind "reference") - else the CPU oracle port
          (kind "port") - fwd+loss+bwd+AdamW on a bounded sample of the same workload, all host threads (rank 0, N=1).
  gpu_eager_baseline : the UNMODIFIED reference modules + torch.optim.AdamW in PyTorch eager (cuBLASLt + SDPA) on the
          same B200 at the same batch, fp32 / bf16 autocast / the trainer's own fp16 autocast + GradScaler recipe
          (train/main.py:75-112); outside the timed region, N=1 only.  This is the bar SURVEY.md names.
Reference arm (--impl reference): the reference's own CPU path for this step: the unmodified reference modules from
/root/reference or, on the GPU box, from the oracle/_ref archive packed from it (kind "reference"); the oracle port
(kind "port") only if neither exists.  Every host thread; each step is a batch-8 sample of the 128-clip step.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

E_LAYERS, D_LAYERS, T, N, D_IN = 6, 6, 64, 32, 4096
PER_GPU_BATCH = 128
CPU_SAMPLE_BATCH = 8
METRIC = "train clips/sec (fwd+bwd) TAN E6D6 T=64"


def train_gflop_per_clip(B):
    """Algorithmic-minimum FLOPs of one training clip: fwd = pre-proj (T+N)*2*D*C +
    video encoder + joint encoder + similarity head (quadratic in B); train = 3x fwd."""
    C, F = 512, 2048
    per_tok = lambda S: 2 * C * 3 * C + 2 * C * C + 4 * C * F + 4 * S * C
    pre = (T + N) * 2 * D_IN * C
    enc = E_LAYERS * T * per_tok(T)
    joint = D_LAYERS * (T + N) * per_tok(T + N)
    sim = (E_LAYERS + D_LAYERS) * T * B * N * 2 * C
    return 3 * (pre + enc + joint + sim) / 1e9


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        busy = [x for x in sm if mx and x > 0.5 * mx] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------ CPU arms
def cpu_step_fn(B, kind_pref="auto"):
    """Returns (kind, cores, callable) timing one fwd+loss+bwd of the CPU path at batch B, fp32."""
    import numpy as np
    import torch
    from exoground_b200.utils import synth
    from exoground_b200.model import TemporalAligner
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    b = synth.tan_batch(B=B, T=T, N=N, d_video=D_IN, d_text=D_IN, zero_pad_text=False)
    args = synth.tan_args(loss_threshold=0.5, use_alignability_head=1)
    shapes = {k: tuple(v.shape) for k, v in TemporalAligner(E_LAYERS, D_LAYERS, use_alignability_head=1,
                                                            video_embed_dim=D_IN, text_embed_dim=D_IN).state_dict().items()}
    sd = synth.fill_state_dict(shapes)
    from oracle import ref_import
    if kind_pref != "port" and ref_import.available():
        r = ref_import.load()
        with ref_import.quiet():
            m = r.tan_model.TemporalAligner(E_LAYERS, D_LAYERS, use_alignability_head=1)
        m.load_state_dict(sd)
        inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
        opt = reference_adamw(m)

        def step():
            opt.zero_grad()
            np.random.seed(0)
            out = m(b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], None)
            with ref_import.quiet():
                L = r.loss.get_loss(inp, b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"],
                                    out, args, None)
            L["loss"].backward()
            opt.step()
            return float(L["loss"])
        return "reference", cores, step
    from oracle import tan_oracle, loss_oracle
    sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    opt = torch.optim.AdamW(list(sd.values()), lr=1e-4, weight_decay=1e-5)

    def step():
        opt.zero_grad()
        np.random.seed(0)
        ps = tan_oracle.draw_pos_starts(T, N)
        out = tan_oracle.tan_forward(sd, E_LAYERS, D_LAYERS, b["video"], b["text"], b["video_padding_mask"],
                                     b["text_padding_mask"], ps, use_alignability_head=1)
        L = loss_oracle.tan_loss(out, b["start"], b["end"], b["text_padding_mask"], T, N, 0.5, 1)
        L["loss"].backward()
        opt.step()
        return float(L["loss"])
    return "port", cores, step


def reference_adamw(model, lr=1e-4, wd=1e-5):
    """train/main.py:350-376,418: AdamW, no weight decay for `.ln_` / `.bias` parameters."""
    import torch
    nd, dc = [], []
    for n, p in model.named_parameters():
        (nd if any(k in n for k in (".ln_", ".bias", ".logit_scale", ".entropy_scale")) else dc).append(p)
    return torch.optim.AdamW([{"params": nd, "lr": lr, "weight_decay": 0.0},
                              {"params": dc, "lr": lr, "weight_decay": wd}], lr=lr, weight_decay=wd)


def gpu_eager_baseline(B, dev, steps=5, warmup=2):
    """The bar SURVEY.md §0/§8(d) names: the reference's own modules (model/tan_model.py, train/loss.py) with
    torch.optim.AdamW in PyTorch eager on THIS GPU, same batch / shapes / loss configuration as the timed step.
    Modes: fp32; bf16 autocast; fp16 autocast + GradScaler (what train/main.py:75-112 actually runs)."""
    import numpy as np
    import torch
    from exoground_b200.utils import synth
    from oracle import ref_import
    if not ref_import.available():
        return {"unavailable": "oracle/_ref archive missing (python -m oracle.build_ref)"}
    r = ref_import.load()
    b = synth.tan_batch(B=B, T=T, N=N, d_video=D_IN, d_text=D_IN, zero_pad_text=False)
    bd = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in b.items()}
    inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
    args = synth.tan_args(loss_threshold=0.5, use_alignability_head=1)
    out = {"what": f"UNMODIFIED reference TemporalAligner E{E_LAYERS}D{D_LAYERS} + train/loss.py get_loss + backward + "
                   f"torch.optim.AdamW, PyTorch eager {torch.__version__}, batch {B}, T={T}, N={N}, D_in={D_IN}, "
                   f"{steps} timed + {warmup} warm-up steps, CUDA events", "unit": "clips/s",
           "source": ref_import.kind()}
    for mode in ("fp32", "bf16_autocast", "fp16_autocast_gradscaler"):
        try:
            torch.manual_seed(888)
            with ref_import.quiet():
                m = r.tan_model.TemporalAligner(E_LAYERS, D_LAYERS, use_alignability_head=1)
            if D_IN != 4096:
                m.video_pre_proj = torch.nn.Linear(D_IN, 512, bias=False)
                m.text_pre_proj = torch.nn.Linear(D_IN, 512, bias=False)
            m = m.to(dev).train()
            opt = reference_adamw(m)
            scaler = torch.amp.GradScaler("cuda", enabled=(mode == "fp16_autocast_gradscaler"))
            ac = {"fp32": None, "bf16_autocast": torch.bfloat16, "fp16_autocast_gradscaler": torch.float16}[mode]

            def step():
                np.random.seed(0)
                with torch.autocast("cuda", dtype=ac, enabled=ac is not None), ref_import.quiet():
                    o = m(bd["video"], bd["text"], bd["video_padding_mask"], bd["text_padding_mask"], None)
                    L = r.loss.get_loss(inp, bd["video"], bd["text"], bd["video_padding_mask"],
                                        bd["text_padding_mask"], o, args, None)
                scaler.scale(L["loss"]).backward()
                scaler.step(opt)
                scaler.update()
                opt.zero_grad()
                return L["loss"]
            for _ in range(warmup):
                step()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                last = step()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            out[mode] = {"value": round(B / (ms * 1e-3), 1), "ms_per_step": round(ms, 3),
                         "last_loss": float(last), "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 2)}
            del m, opt
            torch.cuda.empty_cache()
            torch.cuda.reset_peak_memory_stats()
        except Exception as ex:          # keep the bench line: a failing mode is reported, not fatal
            out[mode] = {"error": f"{type(ex).__name__}: {str(ex)[:200]}"}
            torch.cuda.empty_cache()
    return out


def time_cpu(B, steps, warmup, kind_pref="auto"):
    kind, cores, step = cpu_step_fn(B, kind_pref)
    for _ in range(warmup):
        step()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        step()
        ts.append(time.perf_counter() - t0)
    sec = sum(ts) / len(ts)
    return {"value": B / sec, "unit": "clips/s", "cores": cores, "kind": kind,
            "sample": f"TAN E6D6 fwd + loss + bwd + AdamW, fp32, batch {B} of the {PER_GPU_BATCH}-clip step (T={T}, N={N}, "
                      f"D_in={D_IN}; negatives from the {B} sampled clips), {steps} timed + {warmup} warm-up steps, "
                      f"torch CPU {cores} threads",
            "ms_per_step": sec * 1e3}


def workload_config(n_gpus, xneg=False, batch=0, overlap=False, chunks=1, wire="fp32", host_dtype="bf16", sharded=True):
    global PER_GPU_BATCH
    if batch:
        PER_GPU_BATCH = batch
    cfg = _workload_config(n_gpus)
    cfg["host_features"] = (f"{host_dtype}, pinned, copied every step on a copy stream while the previous step computes"
                            + (" (feature store pre-cast to bf16 offline: same values the fp32 path's first cast kernel "
                               "produces)" if host_dtype == "bf16" else ""))
    if n_gpus > 1:
        cfg["grad_allreduce"] = ("NCCL, per-block buckets (~25 MB) issued asynchronously during backward, inside the "
                                 "step's CUDA graph" if overlap else
                                 f"NCCL reduce-scatter of the {wire} gradients -> AdamW on the owned 1/{n_gpus} of the flat "
                                 f"state -> all-gather of the updated bf16 weights (sharded optimizer state; same "
                                 f"arithmetic as all-reduce + replicated AdamW)" if sharded else
                                 f"NCCL, {wire} gradients after backward in {chunks + 1} parameter-aligned pieces; the "
                                 f"AdamW of piece i is issued when piece i has arrived (under the all-reduce of piece i+1)")
    if xneg and n_gpus > 1:
        cfg["parallelism"] = f"dp{n_gpus} + cross-GPU negatives (text features all-gathered; NCE over the global batch)"
    return cfg


def _workload_config(n_gpus):
    return {"workload": f"TAN E6D6 full train step (fwd + MIL-NCE/alignability loss + bwd + AdamW), T={T}, "
                        f"N={N} sentences, D_in={D_IN}, batch {PER_GPU_BATCH}/GPU (BASELINE.json configs[1]); the CPU "
                        f"reference arm (--impl reference) times a batch-{CPU_SAMPLE_BATCH} sample of this step per "
                        f"step, fp32",
            "global_batch": PER_GPU_BATCH * n_gpus, "per_gpu_batch": PER_GPU_BATCH, "seq_len": T,
            "parallelism": f"dp{n_gpus}", "loss": "loss_threshold=0.5, use_alignability_head=1",
            "l2": "no flush: per-step working set (2 x 403 MB bf16 logits + saved activations, > 3 GB) exceeds the 126 MB L2"}


def nccl_options():
    """Communicator options (experiments with the overlapped gradient all-reduce): EXG_NCCL_HIPRI=1 runs NCCL's
    kernels on a high-priority stream, EXG_NCCL_MAX_CTAS=k caps their CTA count.  Default: NCCL's own choices, which
    measured best (profiles/allreduce_overlap_r01.txt)."""
    import torch.distributed as dist
    opts = dist.ProcessGroupNCCL.Options()
    opts.is_high_priority_stream = os.environ.get("EXG_NCCL_HIPRI", "0") != "0"
    max_ctas = int(os.environ.get("EXG_NCCL_MAX_CTAS", "0"))
    if max_ctas > 0:
        opts.config.max_ctas = max_ctas
    return opts


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = time_cpu(CPU_SAMPLE_BATCH, max(1, a.steps), max(1, min(a.warmup, 2)))
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": "clips/s", "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(a.gpus),
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ------------------------------------------------------------------------------------------ GPU arm
def run_ours(a):
    import numpy as np
    import torch
    import torch.distributed as dist
    from exoground_b200 import ops
    from exoground_b200.ops import lib
    from exoground_b200.model import TemporalAligner
    from exoground_b200.train.step import TanTrainStep
    from exoground_b200.utils import synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (exoground_b200 has no CPU fallback; use --impl reference)")
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    # the rank's thread (and the first touch of its pinned staging pool) on the GPU's own NUMA node
    from exoground_b200.utils import affinity
    a.placement = "unchanged: EXG_NO_BIND" if os.environ.get("EXG_NO_BIND") else affinity.bind_to_gpu(local)
    pg = None
    if world > 1:
        # NCCL writes its version / debug lines to stdout by default; stdout carries the ONE JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device(dev), pg_options=nccl_options())
        pg = dist.group.WORLD
    assert ops.get().name == "cuda"
    B = a.batch or PER_GPU_BATCH

    torch.manual_seed(888)
    np.random.seed(888 + rank)
    model = TemporalAligner(E_LAYERS, D_LAYERS, use_alignability_head=1, video_embed_dim=D_IN,
                            text_embed_dim=D_IN, compute_dtype=torch.bfloat16)
    args = synth.tan_args(loss_threshold=0.5, use_alignability_head=1)
    tr = TanTrainStep(model, args, B, T, N, D_IN, D_IN, device=dev, use_graph=not a.no_graph, process_group=pg,
                      cross_gpu_negatives=a.xneg, overlap_allreduce=a.overlap,
                      comm_sms=int(os.environ.get("EXG_COMM_SMS", "0")), branch_streams=not a.single_stream,
                      early_adamw=a.early_adamw, comm_chunks=a.comm_chunks,
                      comm_dtype=torch.bfloat16 if a.grad_comm == "bf16" else None,
                      host_dtype=torch.bfloat16 if a.host_dtype == "bf16" else torch.float32,
                      shard_optimizer=not a.replicated_optimizer)
    # padded sentence rows hold N(0,1) noise, not zeros: with the reference's default init (LayerNorm bias 0) an
    # all-zero padded row normalises to 0/0 and the REFERENCE's own gradients turn NaN (tan_model.py:110-111)
    batch = synth.tan_batch(B=B, T=T, N=N, d_video=D_IN, d_text=D_IN, seed=888 + rank, zero_pad_text=False)
    if a.host_dtype == "bf16":                     # a feature store kept in half precision (converted offline)
        batch["video"], batch["text"] = batch["video"].bfloat16(), batch["text"].bfloat16()
    for k in ("video", "text", "video_padding_mask", "text_padding_mask"):
        batch[k] = batch[k].pin_memory()
    h2d = tr.load_batch(batch)
    c0 = lib.launch_count()
    tr.prepare()
    torch.cuda.synchronize()

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- launches per step: one eager (un-captured) step goes through every C-ABI entry
    # ---- device-resident timing
    for _ in range(a.warmup):
        tr.run()
    sync_all()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        tr.run()
    e1.record()
    sync_all()
    ms = e0.elapsed_time(e1) / a.steps
    # ---- end-to-end timing through the public step() call
    # the input pipeline is double-buffered: each step() also starts the H2D copy of the NEXT step's batch on the
    # copy stream (same synthetic batch object here), so every timed step contains one full H2D copy + loss read
    # and the loss read-back is pipelined by one step (step_async: every step's loss is copied to pinned host memory and
    # read by the host after the NEXT step has been launched; the last one is read inside the timed region too), so the
    # GPU does not idle through the host's launch work.  --sync-loss restores the blocking step() per iteration.
    for _ in range(min(a.warmup, 3)):
        tr.step(batch, next_batch=batch)
    sync_all()
    t0 = time.perf_counter()
    e0.record()
    last = 0.0
    if a.sync_loss:
        for _ in range(a.steps):
            last = tr.step(batch, next_batch=batch)
    else:
        pending = None
        for _ in range(a.steps):
            h = tr.step_async(batch, next_batch=batch)
            if pending is not None:
                last = pending.result()
            pending = h
        last = pending.result()
    e1.record()
    sync_all()
    ms_e2e = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / a.steps
    clocks = sampler.stop() if rank == 0 else None
    # ---- sustained leg: the same device-resident step back to back for >= 3 s (is the short timed region a burst?)
    t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])
    sustained = None
    if a.sustain_s > 0:
        n_sus = max(a.steps, int(a.sustain_s * 1e3 / ms) + 1)      # from the max-over-ranks time: same count on every rank
        sync_all()
        s2 = ClockSampler(local)
        if rank == 0:
            s2.start()
        e0.record()
        for _ in range(n_sus):
            tr.run()
        e1.record()
        sync_all()
        ms_sus = e0.elapsed_time(e1) / n_sus
        ts = torch.tensor([ms_sus], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ts, op=dist.ReduceOp.MAX)
        ms_sus = float(ts[0])
        sustained = {"value": round(B * world / (ms_sus * 1e-3), 1), "unit": "clips/s", "ms_per_step": round(ms_sus, 4),
                     "steps": n_sus, "seconds": round(ms_sus * n_sus * 1e-3, 2),
                     "clocks": s2.stop() if rank == 0 else None}

    # ---- one profiled eager step: per-kernel-class CUDA-event timings + launch count
    # (single stream for this one step: with the two branches on concurrent streams the per-class event times overlap)
    graph, tr.graph = tr.graph, None
    branch, model._branch_stream = getattr(model, "_branch_stream", None), None
    tr.run(); torch.cuda.synchronize()
    c1 = lib.launch_count()
    lib.profile_begin()
    tr.run()
    prof = lib.profile_end()
    launches_per_step = lib.launch_count() - c1
    tr.graph = graph
    model._branch_stream = branch

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        # the profiled step below is ONE eager step (~10 ms): its kernels run at burst clocks, so the denominator is the
        # burst cuBLAS figure; the sustained figure is reported beside it (and is the one for model_tflops of the
        # `sustained` leg)
        tf_peak = peaks.get("bf16_tflops", 1590.0)
        tf_peak_sus = peaks.get("bf16_tflops_sustained", 1400.0)
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        src = "measured (MEASURED_PEAKS.json, burst)" if peaks else "fallback (B200_PROFILING.md)"
        breakdown = {}
        tot = sum(v["ms"] for v in prof.values())
        for k, v in prof.items():
            tensor = k.startswith("gemm") or k.startswith("attn")
            rate = v["work"] / (v["ms"] * 1e-3) / (1e12 if tensor else 1e9) if v["ms"] > 0 else 0.0
            breakdown[k] = {"ms": round(v["ms"], 4), "launches": v["launches"], "share": round(v["ms"] / tot, 4),
                            ("tflops" if tensor else "gbs"): round(rate, 1)}
        top = max(prof, key=lambda k: prof[k]["ms"])
        tv = prof[top]
        tensor = top.startswith("gemm") or top.startswith("attn")
        ach = tv["work"] / (tv["ms"] * 1e-3) / (1e12 if tensor else 1e9)
        traffic, traffic_src = None, None
        try:      # DRAM bytes per launch of this kernel class from the committed `ncu --set full` capture
            nt = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic_r02.json")))
            traffic = round(nt[top]["dram_bytes_per_launch"])
            traffic_src = nt["source"]
        except Exception:
            pass
        roof = {"kernel": top, "bound": "tensor" if tensor else "hbm", "achieved": round(ach, 2),
                "peak": tf_peak if tensor else hbm_peak, "unit": "TFLOP/s" if tensor else "GB/s",
                "frac": round(ach / (tf_peak if tensor else hbm_peak), 4),
                "frac_of_sustained_peak": round(ach / tf_peak_sus, 4) if tensor else None,
                "peak_sustained": tf_peak_sus if tensor else None, "traffic": traffic,
                "traffic_source": traffic_src, "peak_source": src,
                "launches": tv["launches"], "avg_launch_ms": round(tv["ms"] / tv["launches"], 5),
                "how": "sum of algorithmic work / sum of CUDA-event durations of every launch of this kernel "
                       "class in one profiled eager, single-stream step"}
        gb = B * world
        value = gb / (ms * 1e-3)
        gf = train_gflop_per_clip(B)
        line = {"metric": METRIC, "value": round(value, 1), "unit": "clips/s", "n_gpus": world, "steps": a.steps,
                "warmup": a.warmup, "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": dict(workload_config(world, a.xneg, a.batch, a.overlap, a.comm_chunks, a.grad_comm, a.host_dtype, not a.replicated_optimizer),
                               host_placement=a.placement),
                "e2e": {"value": round(gb / (ms_e2e * 1e-3), 1), "unit": "clips/s", "ms_per_step": round(ms_e2e, 4),
                        "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4, "last_loss": last,
                        "loss_readback": "blocking, every step" if a.sync_loss else
                                         "every step's loss, read one step late (TanTrainStep.step_async)"},
                "gpu_launches": int(launches_per_step * a.steps),
                "launches_per_step": int(launches_per_step), "cuda_graph": not a.no_graph,
                "model_tflops": round(value / world * gf / 1e3, 1),
                "model_tflops_frac_of_peak": round(value / world * gf / 1e3 / tf_peak, 4),
                "model_tflops_frac_of_sustained_peak": round(value / world * gf / 1e3 / tf_peak_sus, 4),
                "train_gflop_per_clip": round(gf, 2),
                "roofline": roof, "breakdown": breakdown, "clocks": clocks}
        if sustained is not None:
            sustained["model_tflops"] = round(sustained["value"] / world * gf / 1e3, 1)
            sustained["model_tflops_frac_of_sustained_peak"] = round(sustained["model_tflops"] / tf_peak_sus, 4)
            line["sustained"] = sustained
        if world == 1 and not a.no_eager:
            del tr, graph
            torch.cuda.empty_cache()
            line["gpu_eager_baseline"] = gpu_eager_baseline(B, dev)
        if world == 1 and not a.no_cpu:
            r = time_cpu(CPU_SAMPLE_BATCH, 3, 1)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        emit(line)
    if world > 1:
        dist.barrier()
        graph = None
        tr.close()              # graphs that captured NCCL kernels must go before their communicator
        dist.destroy_process_group()


_JSON_FD = None


def claim_stdout():
    """stdout carries exactly ONE JSON line: libraries that print there (NCCL's version banner) are sent to stderr
    for the life of the process, and the line is written to the original descriptor at the end."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global T, D_IN
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graph", action="store_true", help="run the step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-eager", action="store_true", help="skip the gpu_eager_baseline leg (reference in PyTorch eager)")
    ap.add_argument("--sustain-s", type=float, default=3.0,
                    help="length of the extra `sustained` leg in seconds (0 = skip)")
    ap.add_argument("--single-stream", action="store_true",
                    help="keep the joint-encoder branch on the main stream (default: its own stream, concurrent with "
                         "the video-encoder branch)")
    ap.add_argument("--early-adamw", action="store_true",
                    help="N=1: launch each transformer stack's AdamW from inside backward as soon as the stack is "
                         "differentiated (default: the whole AdamW after backward; measured: no gain)")
    ap.add_argument("--overlap", action="store_true",
                    help="N>1: per-block gradient buckets all-reduced asynchronously during backward instead of one "
                         "all-reduce after it (measured: no gain, see DESIGN.md §6)")
    ap.add_argument("--comm-chunks", type=int, default=1,
                    help="N>1: chunks of the gradient all-reduce (AdamW of chunk i under the all-reduce of chunk i+1)")
    ap.add_argument("--sync-loss", action="store_true",
                    help="e2e leg: blocking TanTrainStep.step() (loss.item() every iteration) instead of the one-step "
                         "pipelined read-back of step_async()")
    ap.add_argument("--replicated-optimizer", action="store_true",
                    help="N>1: all-reduce + every rank runs the whole AdamW (DistributedDataParallel's scheme) instead of "
                         "reduce-scatter -> AdamW on the owned shard -> all-gather of the bf16 weights")
    ap.add_argument("--grad-comm", default="fp32", choices=["fp32", "bf16"],
                    help="N>1: dtype of the gradients on the wire (fp32 = DistributedDataParallel's; bf16 = its "
                         "bf16_compress_hook)")
    ap.add_argument("--host-dtype", default="bf16", choices=["fp32", "bf16"],
                    help="dtype of the pinned host features: bf16 = a feature store pre-cast offline (bit-identical to the "
                         "on-device cast the fp32 path performs as its first kernel; half the pinned pool / PCIe bytes - "
                         "8 ranks reading fp32 features from one host measured 0.75 end-to-end scaling, bf16 0.91); "
                         "fp32 = what the reference's loaders hand over")
    ap.add_argument("--xneg", action="store_true",
                    help="N > 1: cross-GPU negatives (text features all-gathered, NCE over the global batch; "
                         "train/xneg.py) instead of the DDP-equivalent local negatives")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch override (default: BASELINE configs[1], 128)")
    ap.add_argument("--seq", type=int, default=0, help="T override (BASELINE configs[3]: 256); default 64")
    ap.add_argument("--din", type=int, default=0, help="feature width override (BASELINE configs[3]: 768); default 4096")
    a = ap.parse_args()
    if a.seq:
        T = a.seq
    if a.din:
        D_IN = a.din
    a.warmup = max(a.warmup, 3) if a.impl == "ours" else a.warmup
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
