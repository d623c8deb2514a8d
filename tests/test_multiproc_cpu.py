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
        st = TanTrainStep(m, args, B, T, N, DIN, DIN, process_group=dist.group.WORLD, **kw)
        st.load_batch(b, non_blocking=False)
        st.prepare()
        loss0 = float(st.run())
        fp = m.flat_params()
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
