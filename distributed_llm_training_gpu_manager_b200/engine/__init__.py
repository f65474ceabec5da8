"""Native ZeRO training engine (the part the reference delegates to DeepSpeed)."""
from .config import EngineConfig, load_config
from .zero import ZeroEngine, initialize

__all__ = ["EngineConfig", "load_config", "ZeroEngine", "initialize"]
