"""exoground_b200 — B200-native (sm_100a) implementation of the ExoGround / TAN alignment hot path.

Host code is Python/PyTorch; all arithmetic runs in hand-written CUDA behind the C ABI declared in
include/exg.h (libexg.so).  There is no CPU fallback: ops raise if the library is missing.
"""
__version__ = "0.1.0"
