from .collectives import DistCtx  # noqa: F401
from .torch_engine import TorchEngine  # noqa: F401
