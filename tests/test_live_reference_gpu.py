"""`-m gpu` parity tests against the LIVE, UNMODIFIED reference on the same B200.

The reference modules come from oracle/_ref (the archive oracle/build_ref.py packs from /root/reference in the build
container; it travels to the GPU box like libexg.so) and run in fp32 on the GPU; the drop-in runs on the CUDA
kernels through the C ABI in fp32 (SIMT parity kernels: 1e-4 / bit-exact argmax) and bf16 (tcgen05 kernels: 2e-2).
Sizes are the ones BASELINE.json names:
  cfg-2  TAN E6D6, B=128, T=64, N=32, D_in=4096, loss_threshold=0.5 + alignability head  (the benched step)
  cfg-3  ExoGroundingTransformer E6D6 + decoder, T=1024, N=64, 768-d video / 4096-d text (reaches attention_flash.cu)
  cfg-4  TAN E6D6, T=256, N=32, D_in=768 (per-GPU shape of the cross-GPU-negatives config)
  cfg-5  inference sweep T=64 -> 4096 with interpolate_from=64: argmax over time bit-exact in fp32
plus the ragged / degenerate batches of tests/test_edge_cases_cpu.py on the real kernels, pos_enc='sine' and
ViewInvariantEncoder.
"""
import numpy as np
import pytest
import torch

from exoground_b200 import ops
from oracle import ref_import
from tests import live_ref
from tests.test_edge_cases_cpu import CASES, EXO_CASES

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_import.available(),
                                 reason="oracle/_ref archive missing (python -m oracle.build_ref in the build container)")]

DT = [torch.float32, torch.bfloat16]
IDS = ["fp32", "bf16"]


@pytest.fixture(autouse=True)
def _cuda_backend():
    assert torch.cuda.is_available()
    assert ops.get().name == "cuda"
    yield
    torch.cuda.empty_cache()


@pytest.mark.parametrize("dt", DT, ids=IDS)
def test_cfg2_headline_step_vs_live_reference(dt):
    """BASELINE configs[1]: the exact configuration bench.py times."""
    rep = live_ref.run_tan_live("cuda", dt, 6, 6, 128, 64, 32, 4096, 0.5, 1, what=f"cfg2/{dt}")
    print("cfg-2", dt, rep)


@pytest.mark.parametrize("dt", DT, ids=IDS)
def test_cfg2_smooth_loss_vs_live_reference(dt):
    rep = live_ref.run_tan_live("cuda", dt, 6, 6, 128, 64, 32, 4096, 0.0, 0, what=f"cfg2-smooth/{dt}")
    print("cfg-2 smooth", dt, rep)


@pytest.mark.parametrize("dt", DT, ids=IDS)
def test_cfg4_t256_d768_vs_live_reference(dt):
    """BASELINE configs[3] per-GPU shape (Mode B itself: tests/test_xneg_gpu.py, needs 2 GPUs)."""
    rep = live_ref.run_tan_live("cuda", dt, 6, 6, 16, 256, 32, 768, 0.5, 1, what=f"cfg4/{dt}")
    print("cfg-4", dt, rep)


@pytest.mark.parametrize("dt", DT, ids=IDS)
def test_cfg3_exoground_e6d6_t1024_vs_live_reference(dt):
    """BASELINE configs[2]: ExoGround E6D6 with decoder at T=1024, N=64 - the long-sequence attention kernels
    (attention_flash.cu in bf16, the SIMT flash path in fp32) reached through the model."""
    # random_pos_start=0 (SURVEY cfg-3): the table has 1024 rows, so at T=1024 any non-zero random offset makes the
    # REFERENCE itself fail (exo_ground_model.py:310 adds a shorter slice)
    rep = live_ref.run_exo_live("cuda", dt, 6, 6, 2, 1024, 64, 768, 4096, distill=False, what=f"cfg3/{dt}",
                                grad_tol_mult=2.0, random_pos_start=0)
    print("cfg-3", dt, rep)


def test_cfg3_exoground_distill_t64_vs_live_reference():
    """The reference-default ExoGround shape (T=64, N=64, 4096-d) with the ego-distillation InfoNCE."""
    for dt in DT:
        rep = live_ref.run_exo_live("cuda", dt, 6, 6, 4, 64, 64, 4096, 4096, distill=True, svn=True,
                                    what=f"cfg3-distill/{dt}", grad_tol_mult=2.0)
        print("cfg-3 distill", dt, rep)


@pytest.mark.parametrize("T", [64, 128, 512, 1024, 4096])
def test_cfg5_inference_sweep_interpolate_from(T):
    """BASELINE configs[4]: get_text_visual_sim_joint / _dual with interpolate_from=64 (tan_model.py:151-154,231-277);
    fp32: values 1e-4 and argmax over time bit-exact (top-2 gap printed); bf16: 2e-2."""
    N, d_in = 64, 4096
    r, ref, own32 = live_ref.tan_pair(6, 6, d_in, 0, "cuda", torch.float32, random_pos_start=0)
    g = torch.Generator().manual_seed(T)
    video = torch.randn(1, T, d_in, generator=g).cuda()
    text = torch.randn(1, N, d_in, generator=g).cuda()
    with torch.no_grad():
        want = {"joint": ref.get_text_visual_sim_joint(video, text, interpolate_from=64),
                "dual": ref.get_text_visual_sim_dual(video, text, interpolate_from=64)}
        sim_r = (want["joint"][:, -1] + want["dual"][:, -1]) / 2
        top2 = sim_r.topk(2, dim=1).values
        gap = float((top2[:, 0] - top2[:, 1]).min())
        for dt, own in ((torch.float32, own32), (torch.bfloat16, None)):
            if own is None:
                own = live_ref.tan_pair(6, 6, d_in, 0, "cuda", dt, random_pos_start=0)[2]
            got = {"joint": own.get_text_visual_sim_joint(video, text, interpolate_from=64).float(),
                   "dual": own.get_text_visual_sim_dual(video, text, interpolate_from=64).float()}
            for k in want:
                assert got[k].shape == want[k].shape
                err = float((got[k] - want[k]).abs().max())
                scale = float(want[k].abs().max())
                assert err <= (1e-4 if dt == torch.float32 else 2e-2) * scale + 1e-6, (k, dt, T, err, scale)
            if dt == torch.float32:
                sim_o = (got["joint"][:, -1] + got["dual"][:, -1]) / 2
                assert torch.equal(sim_o.argmax(1), sim_r.argmax(1)), f"T={T}: argmax over time differs (gap {gap:.3e})"
    print(f"cfg-5 T={T}: min top-2 gap {gap:.3e}")


@pytest.mark.parametrize("dt", DT, ids=IDS)
@pytest.mark.parametrize("B,T,N,mv,mt,th,head,seed", CASES)
def test_ragged_and_degenerate_batches_on_cuda(B, T, N, mv, mt, th, head, seed, dt):
    """tests/test_edge_cases_cpu.py's cases on the real kernels: one valid second / sentence, one-clip batch, odd
    sizes, rows / columns without positives."""
    live_ref.run_tan_live("cuda", dt, 2, 3, B, T, N, 24, th, head, seed=seed, np_seed=77,
                          batch_kw=dict(min_video=mv, min_text=mt), push_annotation=(th == 0),
                          what=f"edge{seed}/{dt}")


@pytest.mark.parametrize("dt", DT, ids=IDS)
@pytest.mark.parametrize("B,T,N,mt,cd,svn,only,dec,tpos,seed", EXO_CASES)
def test_exoground_ragged_batches_on_cuda(B, T, N, mt, cd, svn, only, dec, tpos, seed, dt):
    live_ref.run_exo_live("cuda", dt, 2, 2, B, T, N, 24, 40, distill=True, cd=cd, svn=svn, only=only, dec=dec,
                          tpos=tpos, seed=seed, min_text=mt, what=f"exo-edge{seed}/{dt}", grad_tol_mult=2.0)


@pytest.mark.parametrize("dt", DT, ids=IDS)
def test_pos_enc_sine_vs_live_reference(dt):
    """model/tfm_model.py:106-148 (`pos_enc='sine'`): table values and a full forward / loss / backward."""
    from exoground_b200.model.tfm_model import get_position_embedding_sine
    r = ref_import.load()
    want = r.tfm_model.get_position_embedding_sine(512, 4096)
    got = get_position_embedding_sine(512, 4096)
    assert got.shape == want.shape and torch.allclose(got, want, rtol=0, atol=1e-6)
    rep = live_ref.run_tan_live("cuda", dt, 2, 3, 4, 32, 8, 64, 0.0, 0, model_kw=dict(pos_enc="sine"),
                                what=f"sine/{dt}")
    print("sine", dt, rep)


@pytest.mark.parametrize("dt", DT, ids=IDS)
def test_view_invariant_encoder_vs_live_reference(dt):
    """model/vi_encoder.py:8-163: video unimodal encoder + stage mean + 512 -> D_v projection, forward and backward."""
    from exoground_b200.model import ViewInvariantEncoder
    from exoground_b200.utils import synth
    fp32 = dt == torch.float32
    r = ref_import.load()
    Dv, B, T = 768, 3, 64
    kw = dict(num_encoder_layers=3, video_embed_dim=Dv, use_distill_nce_loss=True, random_pos_start=1)
    with ref_import.quiet():
        ref = r.vi_encoder.ViewInvariantEncoder(**kw)
    sd = synth.fill_state_dict({k: v.shape for k, v in ref.state_dict().items()}, seed=13)
    ref.load_state_dict(sd)
    own = ViewInvariantEncoder(compute_dtype=dt, **kw)
    own.load_state_dict(sd)
    ref, own = ref.cuda(), own.cuda()
    g = torch.Generator().manual_seed(3)
    video = torch.randn(B, T, Dv, generator=g).cuda()
    ego = torch.randn(B, T, Dv, generator=g).cuda()
    vmask = torch.zeros(B, T, dtype=torch.bool).cuda()
    vmask[1, 50:] = True
    outs = []
    for mod in (ref, own):
        np.random.seed(9)
        out = mod(video, None, vmask, None, egocentric_video_embed=ego)
        loss = mod.compute_info_nce_loss(out["high_dim_features"], ego) + out["low_dim_features"].square().mean()
        loss.backward()
        outs.append((out, float(loss)))
    (o_r, l_r), (o_o, l_o) = outs
    assert set(o_r) == set(o_o)
    for k in o_r:
        a, c = o_o[k].detach().float(), o_r[k].detach().float()
        err, scale = float((a - c).abs().max()), float(c.abs().max())
        assert a.shape == c.shape and err <= (1e-4 if fp32 else 2e-2) * scale + 1e-6, (k, err, scale)
    assert abs(l_o - l_r) <= (1e-4 if fp32 else 2e-2) * max(1.0, abs(l_r)), (l_o, l_r)
    live_ref.check_grads(own, ref, fp32, what=f"vi_encoder/{dt}")
