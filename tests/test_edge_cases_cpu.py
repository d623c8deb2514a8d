"""Edge cases of the TAN / ExoGround path against the UNMODIFIED reference run live: ragged clips down to one valid
second / one valid sentence, a single-clip batch, sizes that are not multiples of any tile, annotations that fall
outside the clip (rows / columns without a positive), thresholded + alignability losses.  Here the drop-in runs on the
op-spec backend (CPU, fp32: host logic); tests/test_live_reference_gpu.py runs the SAME cases (`CASES`, `EXO_CASES`,
same bodies in tests/live_ref.py) through the CUDA kernels in fp32 and bf16 against the reference on the GPU."""
import pytest
import torch

from exoground_b200 import ops
from oracle import ref_import
from tests import live_ref
from tests.fake_backend import TorchBackend

pytestmark = pytest.mark.skipif(not ref_import.available(), reason="reference (tree or oracle/_ref archive) not present")


@pytest.fixture(autouse=True)
def fake_backend():
    prev = ops.set_backend(TorchBackend())
    yield
    ops.set_backend(prev)


CASES = [
    # B, T, N, min_video, min_text, threshold, head, seed
    (4, 12, 6, 1, 1, 0.0, 0, 1),         # clips with a single valid second / sentence
    (1, 9, 5, 9, 5, 0.0, 0, 2),          # one clip: every negative comes from the clip itself
    (3, 20, 7, 2, 2, 0.5, 1, 3),         # odd sizes + quantile threshold + alignability BCE
    (5, 16, 8, 3, 1, 0.3, 0, 4),
]


@pytest.mark.parametrize("B,T,N,mv,mt,th,head,seed", CASES)
def test_ragged_and_degenerate_batches_match_reference(B, T, N, mv, mt, th, head, seed):
    # (an annotation is pushed past the end of the clip only without the threshold: with it the reference indexes a
    # vector over the columns WITH positives by a mask over the VALID columns and raises IndexError, train/loss.py:299)
    live_ref.run_tan_live("cpu", torch.float32, 2, 3, B, T, N, 24, th, head, seed=seed, np_seed=77,
                          batch_kw=dict(min_video=mv, min_text=mt), push_annotation=(th == 0), what=f"edge{seed}")


EXO_CASES = [
    # B, T, N, min_text, use_center_duration, same_view_negative, only_same, decoder, text_pos, seed
    (3, 10, 5, 1, True, True, False, True, 0, 1),       # clips down to ONE narration, odd T / N
    (1, 7, 3, 3, False, False, False, True, 1, 2),      # a single clip, start / end parametrisation, text pos-enc
    (2, 12, 6, 2, True, True, True, False, 0, 3),       # no decoder: head on every joint stage; only-same-view negatives
]


@pytest.mark.parametrize("B,T,N,mt,cd,svn,only,dec,tpos,seed", EXO_CASES)
def test_exoground_ragged_batches_match_reference(B, T, N, mt, cd, svn, only, dec, tpos, seed):
    """ExoGroundingTransformer + loss_egoexo4d.get_loss + backward against the reference run live on ragged batches."""
    live_ref.run_exo_live("cpu", torch.float32, 2, 2, B, T, N, 24, 40, distill=True, cd=cd, svn=svn, only=only,
                          dec=dec, tpos=tpos, seed=seed, min_text=mt, what=f"exo-edge{seed}", grad_tol_mult=2.0)
