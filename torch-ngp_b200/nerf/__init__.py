from .trainer import NGPTrainer  # noqa: F401
from .network import NeRFNetwork, trunc_exp  # noqa: F401
from .sdf_trainer import SDFTrainer  # noqa: F401
from .rays import get_rays  # noqa: F401
