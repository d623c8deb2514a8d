"""Op layer.  `get()` returns the active backend; the only shipped backend is the CUDA one
(exoground_b200/ops/backend.py -> libexg.so).  `set_backend` exists so that tests can inject a
torch re-statement of the op specs (tests/fake_backend.py) to validate host logic on a box without
a GPU; nothing inside this package ever installs a non-CUDA backend."""
from . import lib
from .lib import (F32, BF16, EPI_NONE, EPI_QGELU, EPI_RESID, EPI_QGELU_BWD, EPI_RELU, EPI_RELU_BWD,
                  GEMM_AUTO, GEMM_SIMT, GEMM_TC, GEMM_TC2, ExgError)

_backend = None


def get():
    global _backend
    if _backend is None:
        from .backend import CudaBackend
        _backend = CudaBackend()          # raises ExgError if libexg.so is missing
    return _backend


def set_backend(b):
    """Test hook.  Returns the previous backend."""
    global _backend
    prev, _backend = _backend, b
    return prev
