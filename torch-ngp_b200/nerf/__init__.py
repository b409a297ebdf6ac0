from .trainer import NGPTrainer  # noqa: F401
from .network import NeRFNetwork, trunc_exp  # noqa: F401
