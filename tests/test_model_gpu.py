"""GPU parity tests proper: the drop-in TemporalAligner + get_loss + backward, running on the CUDA
kernels through the C ABI, against golden fixtures produced by the unmodified reference."""
import numpy as np
import pytest
import torch

from exoground_b200 import ops
from tests.tan_cases import load_golden, run_tan_case

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _cuda_backend():
    assert torch.cuda.is_available()
    assert ops.get().name == "cuda"


def test_tan_small_fp32(golden_dir):
    run_tan_case(load_golden(golden_dir, "tan_small"), [(0.0, 0, 0), (0.3, 0, 0), (0.0, 1, 1), (0.5, 1, 0)],
                 "cuda", torch.float32)


def test_tan_small_bf16(golden_dir):
    run_tan_case(load_golden(golden_dir, "tan_small"), [(0.0, 0, 0), (0.3, 0, 0), (0.0, 1, 1), (0.5, 1, 0)],
                 "cuda", torch.bfloat16)


def test_tan_cfg1_fp32(golden_dir):
    """BASELINE.json configs[0]: TAN E6D6, B=8, T=64, N=32 (D_in=1024), fp32."""
    run_tan_case(load_golden(golden_dir, "tan_cfg1"), [(0.0, 0, 0), (0.5, 1, 0)], "cuda", torch.float32, sub=(8, 4))


def test_tan_cfg1_bf16(golden_dir):
    run_tan_case(load_golden(golden_dir, "tan_cfg1"), [(0.0, 0, 0), (0.5, 1, 0)], "cuda", torch.bfloat16, sub=(8, 4))


def test_launches_counted():
    before = ops.lib.launch_count()
    x = torch.randn(8, 512, device="cuda")
    ops.get().cast(x, torch.bfloat16)
    assert ops.lib.launch_count() == before + 1


# ---------------------------------------------------------------------- ExoGround family
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_exo_small(golden_dir, dt):
    from tests.exo_cases import run_exo_case
    from tests.tan_cases import load_golden
    run_exo_case(load_golden(golden_dir, "exo_small"), "cuda", dt)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_keystep_small(golden_dir, dt):
    from tests.exo_cases import run_keystep_case
    from tests.tan_cases import load_golden
    run_keystep_case(load_golden(golden_dir, "keystep_small"), load_golden(golden_dir, "exo_small"), "cuda", dt)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_exo_audio_and_pairwise(golden_dir, dt):
    from tests.exo_cases import run_audio_pairwise_case
    from tests.tan_cases import load_golden
    run_audio_pairwise_case(load_golden(golden_dir, "exo_audio"), "cuda", dt)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_vi_mlp_grads(dt):
    from tests.exo_cases import run_vi_mlp_grad_case
    run_vi_mlp_grad_case("cuda", dt)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_tan_self_labelling_agreement(golden_dir, dt):
    from tests.tan_cases import run_agree_case
    run_agree_case(load_golden(golden_dir, "tan_agree"), "cuda", dt)


@pytest.mark.parametrize("dt,tol", [(torch.float32, 2e-4), (torch.bfloat16, 3e-2)])
def test_sliding_window_alignment(golden_dir, dt, tol):
    """eval/eval_zeroshot_align.py::test_alignment_htm on synthetic HTM-Align videos: per-sentence scores within
    tolerance; fp32 Recall / AUC / predicted seconds exact; one batched forward == the window loop."""
    from tests.align_cases import run_align_case
    run_align_case(load_golden(golden_dir, "tan_align"), "cuda", dt, tol)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_clip_retrieval_head(dt):
    """eval/eval_zeroshot_retrieval.py:157-258 (YouCook2 retrieval head): features 1e-4 / 2e-2, fp32 metrics exact."""
    from oracle import ref_import
    if not ref_import.available():
        pytest.skip("oracle/_ref archive missing")
    from tests.retrieval_cases import run_retrieval_case
    run_retrieval_case("cuda", dt)


def test_graphed_similarity_matches_eager_helpers():
    """eval/graphed.py: CUDA-graph replay of get_text_visual_sim_joint / _dual == the eager helpers, on new inputs after
    the capture; shape changes and random offsets are refused."""
    import pytest as _pytest
    from exoground_b200.eval.graphed import GraphedSimilarity
    from exoground_b200.model import TemporalAligner
    torch.manual_seed(3)
    m = TemporalAligner(2, 2, use_alignability_head=1, video_embed_dim=96, text_embed_dim=160, random_pos_start=0,
                        compute_dtype=torch.bfloat16).cuda().eval()
    B, T, N = 5, 128, 12
    gs = GraphedSimilarity(m, B, T, N, 96, 160, interpolate_from=64)
    for seed in (1, 2):
        g = torch.Generator().manual_seed(seed)
        video = torch.randn(B, T, 96, generator=g).cuda()
        text = torch.randn(B, N, 160, generator=g).cuda()
        with torch.no_grad():
            want_j = m.get_text_visual_sim_joint(video, text, interpolate_from=64)
            want_d = m.get_text_visual_sim_dual(video, text, interpolate_from=64)
        got_j, got_d = gs(video, text)
        assert torch.equal(got_j, want_j) and torch.equal(got_d, want_d)
    with _pytest.raises(ValueError):
        gs(video[:, :64], text)
    m.random_pos_start = 1
    with _pytest.raises(ValueError):
        GraphedSimilarity(m, B, T, N, 96, 160)
