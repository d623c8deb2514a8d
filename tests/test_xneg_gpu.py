"""Cross-GPU negatives on real GPUs (needs >= 2; skipped otherwise): two NCCL ranks running TanTrainStep with
cross_gpu_negatives=True (libexg.so kernels, CUDA graph replay included) must reproduce the single-GPU step on the
concatenated batch: loss, summed flat gradient."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.test_multiproc_cpu import _cat_batches, _free_port

pytestmark = pytest.mark.gpu
E, D, B, T, N, DIN = 2, 3, 8, 64, 32, 256


def _worker(rank, world, port, q, dtype_name, use_graph):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = f"cuda:{rank}"
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(dev))
    try:
        from exoground_b200 import ops
        from exoground_b200.model import TemporalAligner
        from exoground_b200.train.step import TanTrainStep
        from exoground_b200.utils import synth
        from tests.shapes import tan_shapes
        assert ops.get().name == "cuda"
        dt = getattr(torch, dtype_name)
        args = synth.tan_args(loss_threshold=0.5, use_alignability_head=1)
        kw = dict(lr=1e-3, wd=1e-2, device=dev, use_graph=use_graph)

        def model():
            m = TemporalAligner(E, D, use_alignability_head=1, video_embed_dim=DIN, text_embed_dim=DIN,
                                random_pos_start=0, compute_dtype=dt)
            m.load_state_dict(synth.fill_state_dict(tan_shapes(E, D, DIN, head=1)))
            return m

        batches = [synth.tan_batch(B=B, T=T, N=N, d_video=DIN, d_text=DIN, seed=300 + r, zero_pad_text=False)
                   for r in range(world)]
        m = model()
        st = TanTrainStep(m, args, B, T, N, DIN, DIN, process_group=dist.group.WORLD, cross_gpu_negatives=True, **kw)
        st.load_batch(batches[rank], non_blocking=False)
        st.prepare()
        loss = float(st.run())
        fp = m.flat_params()
        g_dist = fp.grad[:fp.n_used].clone()

        mr = model()
        sr = TanTrainStep(mr, args, world * B, T, N, DIN, DIN, **kw)
        sr.load_batch(_cat_batches(batches), non_blocking=False)
        sr.prepare()
        loss_ref = float(sr.run())
        fr = mr.flat_params()
        g_ref = fr.grad[:fr.n_used].clone()
        tol_l, tol_g = (1e-4, 2e-3) if dt == torch.float32 else (2e-2, 6e-2)
        assert abs(loss - loss_ref) <= tol_l * max(1.0, abs(loss_ref)), (loss, loss_ref)
        err = float((g_dist - g_ref).norm() / g_ref.norm())
        assert err < tol_g, f"gradient of the global loss: rel-L2 {err:.3e}"
        st.close(); sr.close()
        q.put((rank, "ok", (loss, loss_ref, err)))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "fail", traceback.format_exc() + repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
@pytest.mark.parametrize("dtype_name,use_graph", [("float32", False), ("bfloat16", True)])
def test_two_gpu_cross_negatives(dtype_name, use_graph):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q, dtype_name, use_graph)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=500) for _ in procs]
    for rank, status, info in res:
        assert status == "ok", f"rank {rank}: {info}"
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0, "worker did not shut down cleanly (hung in destroy_process_group?)"


# ------------------------------------------------------------------ sharded optimizer == replicated optimizer (NCCL)
def _shard_worker(rank, world, port, q, clip):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = f"cuda:{rank}"
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(dev))
    try:
        from exoground_b200.model import TemporalAligner
        from exoground_b200.train.step import TanTrainStep
        from exoground_b200.utils import synth
        from tests.shapes import tan_shapes
        args = synth.tan_args(loss_threshold=0.5, use_alignability_head=1)
        batches = [synth.tan_batch(B=B, T=T, N=N, d_video=DIN, d_text=DIN, seed=700 + rank + 10 * i, zero_pad_text=False)
                   for i in range(3)]
        out = []
        lr = 1e-4
        for shard in (False, True):
            m = TemporalAligner(E, D, use_alignability_head=1, video_embed_dim=DIN, text_embed_dim=DIN,
                                random_pos_start=0, compute_dtype=torch.bfloat16)
            m.load_state_dict(synth.fill_state_dict(tan_shapes(E, D, DIN, head=1)))
            st = TanTrainStep(m, args, B, T, N, DIN, DIN, lr=lr, wd=1e-2, device=dev, use_graph=True, clip_grad=clip,
                              process_group=dist.group.WORLD, shard_optimizer=shard)
            st.load_batch(batches[0], non_blocking=False)
            st.prepare()                              # settles the flat layout (parameters without gradients move)
            fp = m.flat_params()
            p_init = fp.flat[:fp.n_used].clone()
            l_first = st.step(batches[0])
            st.sync_full()
            fp = m.flat_params()
            first = (fp.flat[:fp.n_used].clone(), fp.shadow[:fp.n_used].clone(), st.m1[:fp.n_used].clone())
            losses = [l_first] + [st.step(b) for b in batches[1:]]
            st.sync_full()
            sd = st.state_dict()
            out.append((losses, first, fp.shadow[:fp.n_used].clone(), sd, p_init))
            if shard:
                assert st._shard_plan is not None and st._shard_plan.S * world >= fp.n_used
            st.close()
        (l0, f0, s0, sd0, pi), (l1, f1, s1, sd1, _) = out
        # The kernels' split-K / atomic reductions make two runs differ in the last bits, so "same update" is checked on
        # the FIRST step from identical weights: every parameter moved, by the same amount as under the replicated
        # optimizer (AdamW's first step is ~lr * sign(g): a rounding-level gradient difference can only flip entries
        # whose gradient is ~0), with the same first moment; later steps: same losses to bf16 precision.
        assert abs(l0[0] - l1[0]) <= 1e-5 * abs(l0[0]), (l0, l1)
        assert all(abs(a - b) <= 3e-2 * abs(a) for a, b in zip(l0, l1)), (l0, l1)
        moved = (f1[0] - pi).abs() > 0
        assert float(moved.float().mean()) > 0.95, "the sharded step left parameters of some shard untouched"
        d = (f0[0] - f1[0]).abs()
        assert float(d.max()) <= 2.1 * lr and float((d > 1e-6).float().mean()) < 1e-2, (d.max(), (d > 1e-6).float().mean())
        assert float((f0[2] - f1[2]).norm() / f0[2].norm()) < 1e-3
        assert torch.equal(f1[1], f1[0].to(torch.bfloat16))          # the gathered bf16 weights ARE the updated masters
        assert sd0["state"].keys() == sd1["state"].keys()
        assert all(float(v["exp_avg"].abs().max()) > 0 for v in sd1["state"].values())
        both = [torch.empty_like(s1) for _ in range(world)]
        dist.all_gather(both, s1)
        assert torch.equal(both[0], both[1])          # identical bf16 weights on both ranks after three steps
        q.put((rank, "ok", l1))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "fail", traceback.format_exc() + repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
@pytest.mark.parametrize("clip", [0.0, 0.5])
def test_two_gpu_sharded_optimizer_matches_replicated(clip):
    """train/flat.py::sharded_reduce_and_update (reduce-scatter -> AdamW on the owned shard -> all-gather of the bf16
    weights) against the replicated all-reduce + AdamW on two NCCL ranks, CUDA-graph step: the first update (fp32
    masters, bf16 shadow, first moment after sync_full) matches, three steps' losses agree, replicas stay identical."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_shard_worker, args=(r, 2, port, q, clip)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=500) for _ in procs]
    for rank, status, info in res:
        assert status == "ok", f"rank {rank}: {info}"
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
