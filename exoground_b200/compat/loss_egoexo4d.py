"""Shim: `from loss_egoexo4d import get_loss, get_mask_from_time, get_text_pos, visualize, save_features_to_dir`
(train/main_egoexo4d_distributed.py:18) -> the B200 drop-in.  `visualize` / `save_features_to_dir` (matplotlib /
ffmpeg / file I/O, off the hot path; loss_egoexo4d.py:365-566) are forwarded to the reference's own module, loaded
from its file the first time they are called."""
import importlib.machinery
import importlib.util
import os
import sys

from exoground_b200.train.loss_egoexo4d import (get_loss, get_mask_from_time, get_text_pos,  # noqa: F401
                                                get_grounding_loss_reg_head, get_view_invariant_loss)

_ref = None


def _reference_module():
    global _ref
    if _ref is None:
        here = os.path.dirname(os.path.abspath(__file__))
        for p in sys.path:
            if not p or os.path.abspath(p) == here:
                continue
            spec = importlib.machinery.PathFinder.find_spec("loss_egoexo4d", [p])     # directories and zip archives
            if spec is not None and spec.loader is not None:
                _ref = importlib.util.module_from_spec(spec)
                spec.loader.exec_module(_ref)
                break
        else:
            raise ImportError("the reference's train/loss_egoexo4d.py is not on sys.path (needed for visualize / "
                              "save_features_to_dir only)")
    return _ref


def visualize(*args, **kwargs):
    return _reference_module().visualize(*args, **kwargs)


def save_features_to_dir(*args, **kwargs):
    return _reference_module().save_features_to_dir(*args, **kwargs)
