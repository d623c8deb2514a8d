"""Model classes with the reference's names and signatures (model/*.py of arjunrs1/ExoGround)."""
from .tan_model import TemporalAligner, TwinTemporalAligner  # noqa: F401
