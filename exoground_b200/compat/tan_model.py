"""Shim: `from tan_model import TemporalAligner, TwinTemporalAligner` (train/main.py:24) -> the B200 drop-in."""
from exoground_b200.model.tan_model import TemporalAligner, TwinTemporalAligner  # noqa: F401
