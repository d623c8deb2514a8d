"""TEST INFRASTRUCTURE ONLY.  Import the UNMODIFIED reference modules.

Source of the modules, in order:
  1. the reference tree itself (/root/reference, or $EXG_REFERENCE_ROOT) -- the build container;
  2. `oracle/_ref/exoground_reference.pyz`, the archive `oracle/build_ref.py` packs from that tree (git-ignored
     build product that travels to the GPU box with the snapshot) -- imported through zipimport.
Used by oracle/make_golden.py, the CPU tests that drive the live reference, the `-m gpu` live-reference parity
tests and bench.py's reference legs.  Nothing under exoground_b200/ imports this.

Stub modules replace plotting/video packages that the reference imports at module top but never uses in the
math (train/loss.py:4,12; train/loss_egoexo4d.py:4,12-15; utils/utils.py:11,15).
"""
import contextlib
import io
import os
import sys
import types

REF = os.environ.get("EXG_REFERENCE_ROOT", "/root/reference")
ARCHIVE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "exoground_reference.pyz")


def root():
    """Path prefix the reference's packages hang off: the tree, else the packed archive, else None."""
    if os.environ.get("EXG_REFERENCE_FORCE_ARCHIVE") != "1" and os.path.isdir(os.path.join(REF, "model")):
        return REF
    if os.path.exists(ARCHIVE):
        return ARCHIVE
    return None


def kind():
    r = root()
    return None if r is None else ("tree" if r == REF else "archive")


def available():
    return root() is not None


def _stub(name, **attrs):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def _install_stubs():
    try:
        import matplotlib  # noqa: F401
        import matplotlib.pyplot  # noqa: F401
    except Exception:
        mpl = _stub("matplotlib", use=lambda *a, **k: None)
        plt = _stub("matplotlib.pyplot", switch_backend=lambda *a, **k: None)
        mpl.pyplot = plt
    for name in ("ffmpeg", "moviepy", "moviepy.editor", "tensorboardX", "gpustat", "cv2"):
        try:
            __import__(name)
        except Exception:
            _stub(name, VideoFileClip=object, SummaryWriter=object)
    try:
        import torchvision  # noqa: F401
        from torchvision import transforms  # noqa: F401
    except Exception:
        tv = _stub("torchvision")
        tv.transforms = _stub("torchvision.transforms")


def paths():
    r = root()
    if r is None:
        raise RuntimeError(f"reference not found: no tree at {REF} and no archive at {ARCHIVE} "
                           "(python -m oracle.build_ref packs it in the build container)")
    return [os.path.join(r, "model"), r, os.path.join(r, "train")]


def load():
    """Returns a namespace of reference modules: tfm_model, tan_model, exo_ground_model,
    keystep_ground_model, vi_encoder, loss, loss_egoexo4d."""
    for p in paths():
        if p not in sys.path:
            sys.path.insert(0, p)
    _install_stubs()
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
