from .ffmlp import FFMLP, ffmlp_forward, convert_activation, available, _backend  # noqa: F401
