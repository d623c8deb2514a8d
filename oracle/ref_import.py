"""TEST INFRASTRUCTURE ONLY.  Import the UNMODIFIED reference modules from /root/reference.

Only usable in the build container (the GPU box has no /root/reference); used by
oracle/make_golden.py and by CPU tests that are skipped when the tree is absent.
Stub modules replace plotting/video packages that the reference imports at module
top but never uses in the math (train/loss.py:4,12; train/loss_egoexo4d.py:4,12-15;
utils/utils.py:11).
"""
import contextlib
import io
import os
import sys
import types

REF = os.environ.get("EXG_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REF, "model"))


def _stub(name, **attrs):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def load():
    """Returns a namespace of reference modules: tfm_model, tan_model, exo_ground_model,
    keystep_ground_model, vi_encoder, loss, loss_egoexo4d."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REF}")
    for p in (os.path.join(REF, "model"), REF, os.path.join(REF, "train")):
        if p not in sys.path:
            sys.path.insert(0, p)
    try:
        import matplotlib  # noqa: F401
    except Exception:
        mpl = _stub("matplotlib", use=lambda *a, **k: None)
        plt = _stub("matplotlib.pyplot", switch_backend=lambda *a, **k: None)
        mpl.pyplot = plt
    for name in ("ffmpeg", "moviepy", "moviepy.editor", "tensorboardX", "gpustat"):
        try:
            __import__(name)
        except Exception:
            _stub(name, VideoFileClip=object, SummaryWriter=object)
    with contextlib.redirect_stdout(io.StringIO()):
        import tfm_model, tan_model, exo_ground_model, keystep_ground_model, vi_encoder  # noqa: E401
        import loss as loss_tan
        import loss_egoexo4d
    return types.SimpleNamespace(tfm_model=tfm_model, tan_model=tan_model,
                                 exo_ground_model=exo_ground_model,
                                 keystep_ground_model=keystep_ground_model,
                                 vi_encoder=vi_encoder, loss=loss_tan, loss_egoexo4d=loss_egoexo4d)


@contextlib.contextmanager
def quiet():
    """The reference prints shapes on every loss call (train/loss.py:40-41,77)."""
    with contextlib.redirect_stdout(io.StringIO()):
        yield
