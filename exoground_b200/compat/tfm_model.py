"""Shim: `from tfm_model import ...` (model/tan_model.py:9, exo_ground_model.py:9) -> the B200 drop-in's holders."""
from exoground_b200.model.tfm_model import (TemporalEncoder, TemporalDecoder, get_position_embedding_sine,  # noqa: F401
                                            QuickGELU, ResidualAttentionBlock_Step, ResidualDecoderBlock_Step)
