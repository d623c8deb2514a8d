"""N>1 data-parallel logic on CPU: two `gloo` processes run TanTrainStep (op-spec backend, fp32) on different
batches; the all-reduced flat gradient must be the sum of the per-rank gradients, the fused AdamW must apply their
mean (DDP semantics, reference train/main.py:413), and the parameters must stay identical on both ranks."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

E, D, B, T, N, DIN = 1, 1, 3, 8, 4, 32


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make(rank_seed):
    from exoground_b200.model import TemporalAligner
    from exoground_b200.utils import synth
    from tests.shapes import tan_shapes
    m = TemporalAligner(E, D, use_alignability_head=0, video_embed_dim=DIN, text_embed_dim=DIN, random_pos_start=0,
                        compute_dtype=torch.float32)
    m.load_state_dict(synth.fill_state_dict(tan_shapes(E, D, DIN, head=0)))
    b = synth.tan_batch(B=B, T=T, N=N, d_video=DIN, d_text=DIN, seed=100 + rank_seed)
    return m, b


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from exoground_b200 import ops
        from exoground_b200.train.step import TanTrainStep
        from exoground_b200.utils import synth
        from tests.fake_backend import TorchBackend
        ops.set_backend(TorchBackend())
        args = synth.tan_args(loss_threshold=0.0, use_alignability_head=0)
        kw = dict(lr=1e-3, wd=1e-2, device="cpu", use_graph=False)

        # distributed run: this rank's batch only
        m, b = _make(rank)
        # (overlapped all-reduce: one asynchronous bucket per ~2 transformer blocks + the remainder at the end)
        st = TanTrainStep(m, args, B, T, N, DIN, DIN, process_group=dist.group.WORLD, overlap_allreduce=True,
                          bucket_mb=8.0, **kw)
        st.load_batch(b, non_blocking=False)
        st.prepare()
        calls = []
        orig_ar = dist.all_reduce
        dist.all_reduce = lambda t, **k: (calls.append(t.numel()), orig_ar(t, **k))[1]
        loss0 = float(st.run())
        dist.all_reduce = orig_ar
        fp = m.flat_params()
        assert st._hooked and len(fp.layer_ranges) == E + D and len(calls) >= 3, (calls, fp.layer_ranges)
        assert sum(calls) == fp.n_used and calls[-1] == fp.n_bucketed
        summed = fp.grad[:fp.n_used].clone()

        # single-process restatement: both ranks' gradients computed locally, mean applied
        grads = []
        for r in range(world):
            mr, br = _make(r)
            sr = TanTrainStep(mr, args, B, T, N, DIN, DIN, **kw)
            sr.load_batch(br, non_blocking=False)
            sr.prepare()
            fr = mr.flat_params()
            fr.grad.zero_()
            sr._draw_offsets()
            sr._fwd_bwd()
            grads.append(fr.grad[:fr.n_used].clone())
        want_sum = sum(grads)
        assert torch.allclose(summed, want_sum, rtol=1e-5, atol=1e-7), (summed - want_sum).abs().max()
        assert (grads[0] - grads[1]).abs().max() > 1e-6            # the two ranks really saw different data
        # expected parameters after one AdamW step on the mean gradient
        mref, _ = _make(0)
        sref = TanTrainStep(mref, args, B, T, N, DIN, DIN, **kw)
        sref.load_batch(_make(0)[1], non_blocking=False)
        sref.prepare()
        fref = mref.flat_params()
        fref.grad.zero_()
        fref.grad[:fref.n_used] = want_sum / world
        sref._optimizer()
        assert torch.allclose(fp.flat, fref.flat, rtol=1e-5, atol=1e-7), (fp.flat - fref.flat).abs().max()

        # a second step, then the replicas must still agree bit for bit
        st.run()
        mine = fp.flat.clone()
        both = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(both, mine)
        assert torch.equal(both[0], both[1])
        q.put((rank, "ok", loss0))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "fail", traceback.format_exc() + repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_gradient_allreduce_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, status, info in res:
        assert status == "ok", f"rank {rank}: {info}"
    assert abs(res[0][2] - res[1][2]) > 0        # different batches -> different local losses


def _chunked_worker(rank, world, port, q, clip, bf16_wire, shard=False):
    """The default multi-GPU path: chunked all-reduce with the AdamW of chunk i issued as soon as chunk i has arrived."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from exoground_b200 import ops
        from exoground_b200.train.step import TanTrainStep
        from exoground_b200.utils import synth
        from tests.fake_backend import TorchBackend
        ops.set_backend(TorchBackend())
        args = synth.tan_args(loss_threshold=0.0, use_alignability_head=0)
        kw = dict(lr=1e-3, wd=1e-2, device="cpu", use_graph=False, clip_grad=clip)
        m, b = _make(rank)
        st = TanTrainStep(m, args, B, T, N, DIN, DIN, process_group=dist.group.WORLD, comm_chunks=3,
                          comm_dtype=torch.bfloat16 if bf16_wire else None, shard_optimizer=shard, **kw)
        st.load_batch(b, non_blocking=False)
        st.prepare()
        calls = []
        orig_ar = dist.all_reduce
        dist.all_reduce = lambda t, **k: (calls.append((t.numel(), t.dtype)), orig_ar(t, **k))[1]
        loss0 = float(st.run())
        dist.all_reduce = orig_ar
        fp = m.flat_params()
        if shard:
            # reduce-scatter -> AdamW on the owned shard -> all-gather of the weights: the only all-reduce is the tiny
            # vector of per-parameter squared norms; moments outside the shard are untouched (still zero)
            sp = st._shard_plan
            assert sp is not None and sp.S % 64 == 0 and sp.padded >= fp.n_used and sp.lo == rank * sp.S
            assert [n for n, _ in calls] == ([len(st.seg_names)] if clip > 0 else []), calls
            other = slice((1 - rank) * sp.S, (1 - rank) * sp.S + sp.S)
            assert float(st.m1[other].abs().max()) == 0.0 and float(st.m1[sp.lo:sp.lo + sp.S].abs().max()) > 0.0
            st.sync_full()
            assert float(st.m1[other].abs().max()) > 0.0
        else:
            plan = st._chunk_plan
            assert not st._hooked and len(plan) >= 3 and len(calls) == len(plan), (plan, calls)
            assert sum(n for n, _ in calls) == fp.n_used and plan[0][0] == 0 and plan[-1][1] == fp.n_used
            assert all(dt == (torch.bfloat16 if bf16_wire else torch.float32) for _, dt in calls)
            assert any(hi == fp.n_no_decay for _, hi, _ in plan) and all(lo % 64 == 0 for lo, _, _ in plan)
            starts = {off for n, (off, numel, _) in fp.index.items()}
            assert all(lo in starts for lo, _, _ in plan)                 # cuts at parameter boundaries only

        grads = []
        for r in range(world):
            mr, br = _make(r)
            sr = TanTrainStep(mr, args, B, T, N, DIN, DIN, **kw)
            sr.load_batch(br, non_blocking=False)
            sr.prepare()
            fr = mr.flat_params()
            fr.grad.zero_()
            sr._draw_offsets()
            sr._fwd_bwd()
            grads.append(fr.grad[:fr.n_used].clone())
        want_sum = sum(grads)
        mref, _ = _make(0)
        sref = TanTrainStep(mref, args, B, T, N, DIN, DIN, **kw)
        sref.load_batch(_make(0)[1], non_blocking=False)
        sref.prepare()
        fref = mref.flat_params()
        fref.grad.zero_()
        fref.grad[:fref.n_used] = want_sum / world
        sref._optimizer()                                              # per-parameter clip + AdamW on the mean gradient
        if clip > 0:
            assert bool((sref.seg_sumsq.sqrt() > clip).any()) and bool((sref.seg_sumsq.sqrt() < clip).any())
        diff = (fp.flat - fref.flat).abs()
        if bf16_wire:
            # the first AdamW step moves every weight by ~lr * sign(g): bf16 rounding on the wire matters only where
            # g is ~0 (and through the clip factor's 3-digit norm)
            assert float(diff.max()) <= 2.1e-3 and float((diff > 2e-5).float().mean()) < 2e-2, (diff.max(),)
        else:
            assert torch.allclose(fp.flat, fref.flat, rtol=1e-5, atol=1e-7), diff.max()
        st.run()
        st.sync_full()
        mine = fp.flat.clone()
        both = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(both, mine)
        assert torch.equal(both[0], both[1])
        if shard:                                   # the checkpointed optimizer state is complete on every rank
            sd = st.state_dict()
            nz = [float(v["exp_avg"].abs().max()) > 0 for v in sd["state"].values()]
            assert sum(nz) >= 0.9 * len(nz), (sum(nz), len(nz))
        q.put((rank, "ok", loss0))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "fail", traceback.format_exc() + repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("clip,bf16_wire,shard", [(0.0, False, False), (0.05, False, False), (0.05, True, False),
                                                  (0.0, False, True), (0.05, False, True), (0.05, True, True)])
def test_two_rank_chunked_allreduce_pipelined_adamw_gloo(clip, bf16_wire, shard):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_chunked_worker, args=(r, 2, port, q, clip, bf16_wire, shard)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, status, info in res:
        assert status == "ok", f"rank {rank}: {info}"


# ----------------------------------------------------------------------------------------------- cross-GPU negatives
def _cat_batches(bs):
    out = {}
    for k in bs[0]:
        if torch.is_tensor(bs[0][k]):
            out[k] = torch.cat([b[k] for b in bs], 0)
        else:
            out[k] = [x for b in bs for x in b[k]]
    return out


def _xneg_worker(rank, world, port, q, head, th):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from exoground_b200 import ops
        from exoground_b200.model import TemporalAligner
        from exoground_b200.train.step import TanTrainStep
        from exoground_b200.utils import synth
        from tests.fake_backend import TorchBackend
        from tests.shapes import tan_shapes
        ops.set_backend(TorchBackend())
        E2, D2 = 2, 3
        args = synth.tan_args(loss_threshold=th, use_alignability_head=head, use_abs_text_pos=bool(head))
        kw = dict(lr=1e-3, wd=1e-2, device="cpu", use_graph=False)

        def model():
            m = TemporalAligner(E2, D2, use_alignability_head=head, video_embed_dim=DIN, text_embed_dim=DIN,
                                random_pos_start=0, compute_dtype=torch.float32)
            m.load_state_dict(synth.fill_state_dict(tan_shapes(E2, D2, DIN, head=head)))
            return m

        batches = [synth.tan_batch(B=B, T=T, N=N, d_video=DIN, d_text=DIN, seed=300 + r) for r in range(world)]
        for b in batches:       # absolute sentence positions (loss.py:337-347 reads their mean)
            g = torch.Generator().manual_seed(int(b["video"].abs().sum() * 1000) % 10007)
            b["abs_text_pos"] = torch.rand(B, N, 2, generator=g)

        # distributed: this rank's clips, negatives from both ranks
        m = model()
        st = TanTrainStep(m, args, B, T, N, DIN, DIN, process_group=dist.group.WORLD, cross_gpu_negatives=True, **kw)
        st.load_batch(batches[rank], non_blocking=False)
        st.prepare()
        loss = float(st.run())
        fp = m.flat_params()
        g_dist = fp.grad[:fp.n_used].clone()

        # single process on the concatenated batch
        mr = model()
        sr = TanTrainStep(mr, args, world * B, T, N, DIN, DIN, **kw)
        sr.load_batch(_cat_batches(batches), non_blocking=False)
        sr.prepare()
        loss_ref = float(sr.run())
        fr = mr.flat_params()
        g_ref = fr.grad[:fr.n_used].clone()
        assert abs(loss - loss_ref) <= 1e-5 * max(1.0, abs(loss_ref)), (loss, loss_ref)
        err = (g_dist - g_ref).norm() / g_ref.norm()
        assert err < 1e-4, f"gradient of the global loss: rel-L2 {err:.3e}"
        # first Adam step ~ lr * sign(g): elements with |g| near eps amplify fp32 summation-order noise
        assert ((fp.flat - fr.flat).abs() > 2e-4).float().mean() < 1e-3, (fp.flat - fr.flat).abs().max()
        q.put((rank, "ok", loss))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "fail", traceback.format_exc() + repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
@pytest.mark.parametrize("head,th", [(0, 0.0), (1, 0.5)])
def test_two_rank_cross_gpu_negatives_gloo(head, th):
    """Mode B == the single-process step on the concatenated batch: loss, summed gradients, updated parameters."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_xneg_worker, args=(r, 2, port, q, head, th)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=500) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, status, info in res:
        assert status == "ok", f"rank {rank}: {info}"
    assert abs(res[0][2] - res[1][2]) < 1e-6      # both ranks report the same (global) loss


# ----------------------------------------------------------------------------------------------- ExoGround DDP replacement
def _exo_worker(rank, world, port, q, shard=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import numpy as np
        from exoground_b200 import ops
        from exoground_b200.model import ExoGroundingTransformer
        from exoground_b200.train.step_exo import ExoTrainStep
        from exoground_b200.utils import synth
        from tests.fake_backend import TorchBackend
        from tests.shapes import exo_shapes
        ops.set_backend(TorchBackend())
        Dv, Dt, Bx, Tx, Nx = 24, 40, 3, 16, 6
        args = synth.exo_args(use_distill_nce_loss=True, same_view_negative=True)

        def make():
            m = ExoGroundingTransformer(1, 1, video_embed_dim=Dv, text_embed_dim=Dt, use_distill_nce_loss=True,
                                        random_pos_start=0, compute_dtype=torch.float32)
            m.load_state_dict(synth.fill_state_dict(exo_shapes(1, 1, Dv, Dt, distill=True)))
            return m

        batches = [synth.exo_batch(B=Bx, T=Tx, N=Nx, d_video=Dv, d_text=Dt, d_ego=Dv, seed=500 + r) for r in range(world)]
        st = ExoTrainStep(make(), args, lr=1e-3, wd=1e-2, clip_grad=3.0, device="cpu", process_group=dist.group.WORLD,
                          shard_optimizer=shard)
        L = st.step(batches[rank])
        g_dist = st.opt.flat_grad().clone()            # all-reduced (sum); the clip is folded into the AdamW read
        if shard:                                      # reduce-scatter: only the owned shard holds the sum
            sp = st.opt._shard_plan
            both_g = [torch.zeros(sp.S) for _ in range(world)]
            gpad = torch.zeros(sp.padded)
            gpad[:g_dist.numel()] = g_dist
            dist.all_gather(both_g, gpad[sp.lo:sp.lo + sp.S].clone())
            g_dist = torch.cat(both_g)[:g_dist.numel()]
            st.opt.sync_full()

        # single-process restatement: mean of the two ranks' gradients, the reference's per-parameter clip, AdamW
        grads, models = [], []
        for r in range(world):
            sr = ExoTrainStep(make(), args, lr=1e-3, wd=1e-2, device="cpu")
            b = sr.to_device(batches[r])
            for p in sr.model.parameters():
                p.grad = None
            sr._fwd_bwd(b)
            sr.opt.attach()
            sr._fwd_bwd(b)
            grads.append(sr.opt.flat_grad().clone())
            models.append(sr)
        assert (grads[0] - grads[1]).abs().max() > 1e-7
        ref = models[0]
        ref.opt.flat_grad().copy_((grads[0] + grads[1]) / world)
        want = ref.opt.flat_grad().clone()
        assert torch.allclose(g_dist / world, want, rtol=1e-5, atol=1e-7), (g_dist / world - want).abs().max()
        clipped = 0
        for _, p in ref.model.named_parameters():      # utils/train_utils.py:3-13
            if p.grad is not None:
                n = p.grad.norm(2)
                c = 3.0 / (n + 1e-6)
                if c < 1:
                    p.grad.mul_(c)
                    clipped += 1
        assert clipped > 0, "case does not exercise the clip"
        ref.opt.step()
        mine, theirs = st.model.flat_params().flat, ref.model.flat_params().flat
        assert ((mine - theirs).abs() > 2e-4).float().mean() < 1e-3, (mine - theirs).abs().max()
        both = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(both, mine.clone())
        assert torch.equal(both[0], both[1])
        q.put((rank, "ok", float(L["loss"])))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "fail", traceback.format_exc() + repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("shard", [False, True], ids=["replicated", "sharded"])
def test_two_rank_exo_step_gloo(shard):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_exo_worker, args=(r, 2, port, q, shard)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, status, info in res:
        assert status == "ok", f"rank {rank}: {info}"
