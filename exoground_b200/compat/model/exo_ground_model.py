"""Shim: `from model.exo_ground_model import ExoGroundingTransformer` (main_egoexo4d_distributed.py:27)."""
from exoground_b200.model.exo_ground_model import ExoGroundingTransformer  # noqa: F401
