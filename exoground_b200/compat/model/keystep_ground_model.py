"""Shim: `from model.keystep_ground_model import GroundingModel` (main_egoexo4d_distributed.py:28)."""
from exoground_b200.model.keystep_ground_model import GroundingModel  # noqa: F401
