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
