from .agent import build_models  # noqa: F401
from .dreamer_v3 import train  # noqa: F401
