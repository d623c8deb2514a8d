"""Shim: `from model.vi_encoder import ViewInvariantEncoder, ViewInvariantMLP` (main_egoexo4d_distributed.py:29)."""
from exoground_b200.model.vi_encoder import ViewInvariantEncoder, ViewInvariantMLP  # noqa: F401
