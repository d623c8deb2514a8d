from .tan_model import TemporalAligner, TwinTemporalAligner  # noqa: F401
from .exo_ground_model import ExoGroundingTransformer  # noqa: F401
from .keystep_ground_model import GroundingModel  # noqa: F401
from .vi_encoder import ViewInvariantMLP, ViewInvariantEncoder  # noqa: F401
