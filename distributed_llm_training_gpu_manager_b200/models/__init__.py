from .configs import PRESET_MODELS, PRESETS, ModelConfig, get_config


def build_torch_model(cfg: ModelConfig):
    """Reference ``nn.Module`` for a config (generic ZeRO path / oracle)."""
    if cfg.family == "gpt2":
        from .gpt2 import GPT2
        return GPT2(cfg)
    if cfg.family == "llama":
        from .llama import Llama
        return Llama(cfg)
    raise ValueError(f"unknown model family {cfg.family!r}")


__all__ = ["PRESETS", "PRESET_MODELS", "ModelConfig", "get_config", "build_torch_model"]
