"""Shim: `from loss import get_loss, get_mask_from_time, get_text_pos` (train/main.py:16) -> the B200 drop-in."""
from exoground_b200.train.loss import get_loss, get_mask_from_time, get_text_pos, circulant  # noqa: F401
