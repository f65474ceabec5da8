"""Model architecture configs and presets.

The reference ships no model code at all (it launches a user script it does not include,
SURVEY.md §0); BASELINE.json names the architectures the framework must train: tiny GPT-2
(2 layers, d=128), GPT-2 XL 1.5B, Llama-3 8B and Llama-3 70B."""
from __future__ import annotations

from dataclasses import asdict, dataclass
from typing import Dict


@dataclass(frozen=True)
class ModelConfig:
    family: str            # "gpt2" | "llama"
    vocab_size: int
    hidden: int
    n_layers: int
    n_heads: int
    n_kv_heads: int
    ffn: int
    max_seq: int
    rope_theta: float = 500000.0
    norm_eps: float = 1e-5
    tie_embeddings: bool = False

    @property
    def head_dim(self) -> int:
        return self.hidden // self.n_heads

    @property
    def padded_vocab(self) -> int:
        """Vocabulary rounded up to a multiple of 256 (GEMM tile / TMA alignment)."""
        return (self.vocab_size + 255) // 256 * 256

    def num_params(self) -> int:
        d, f, v = self.hidden, self.ffn, self.padded_vocab
        if self.family == "llama":
            kv = self.n_kv_heads * self.head_dim
            per_layer = d * (d + 2 * kv) + d * d + 3 * d * f + 2 * d
            return self.n_layers * per_layer + v * d * (1 if self.tie_embeddings else 2) + d
        per_layer = 4 * d * d + 4 * d + 2 * d * f + f + d + 4 * d
        return self.n_layers * per_layer + v * d + self.max_seq * d + 2 * d

    def flops_per_token(self, seq_len: int, recompute: bool = True) -> float:
        """Training FLOPs per token: 6·N_matmul (+2·N_matmul when activations are recomputed)
        plus causal attention (QKᵀ and PV, forward + 2.5× backward [+ forward recompute])."""
        d, f = self.hidden, self.ffn
        kv = self.n_kv_heads * self.head_dim
        if self.family == "llama":
            mm = self.n_layers * (d * (d + 2 * kv) + d * d + 3 * d * f) + self.padded_vocab * d
        else:
            mm = self.n_layers * (4 * d * d + 2 * d * f) + self.padded_vocab * d
        attn_fwd = self.n_layers * 2 * 2 * seq_len * d / 2  # per token, causal half
        mult_mm = 8 if recompute else 6
        mult_attn = 4.5 if recompute else 3.5
        return mult_mm * mm + mult_attn * attn_fwd

    def to_dict(self) -> Dict:
        return asdict(self)


PRESETS: Dict[str, ModelConfig] = {
    # BASELINE.json config 1
    "gpt2-tiny": ModelConfig("gpt2", 512, 128, 2, 4, 4, 512, 128, tie_embeddings=True),
    "gpt2-small": ModelConfig("gpt2", 50257, 768, 12, 12, 12, 3072, 1024, tie_embeddings=True),
    # BASELINE.json config 2
    "gpt2-xl": ModelConfig("gpt2", 50257, 1600, 48, 25, 25, 6400, 1024, tie_embeddings=True),
    "llama-tiny": ModelConfig("llama", 1024, 256, 2, 2, 1, 512, 256),
    "llama-1b": ModelConfig("llama", 128256, 2048, 16, 16, 4, 8192, 8192),
    # BASELINE.json configs 3/4
    "llama3-8b": ModelConfig("llama", 128256, 4096, 32, 32, 8, 14336, 8192),
    # BASELINE.json config 5
    "llama3-70b": ModelConfig("llama", 128256, 8192, 80, 64, 8, 28672, 8192),
    # the sizes behind the launcher's `7b` / `13b` presets (ref ai_engine/deepspeed_launcher.py:373-392):
    # Llama-2 shapes — multi-head attention (kv heads = heads), head_dim 128, so they run on the same kernels
    "llama2-7b": ModelConfig("llama", 32000, 4096, 32, 32, 32, 11008, 4096, rope_theta=10000.0),
    "llama2-13b": ModelConfig("llama", 32000, 5120, 40, 40, 40, 13824, 4096, rope_theta=10000.0),
}

# launcher preset name -> architecture used when a job is started from that preset without naming a model
PRESET_MODELS: Dict[str, str] = {"7b": "llama2-7b", "13b": "llama2-13b", "70b": "llama3-70b"}


def get_config(name: str, **overrides) -> ModelConfig:
    name = PRESET_MODELS.get(name, name)       # "7b" / "13b" / "70b" (launcher preset names) are accepted as aliases
    if name not in PRESETS:
        raise KeyError(f"unknown model preset {name!r}; available: {sorted(PRESETS)}")
    cfg = PRESETS[name]
    if overrides:
        d = cfg.to_dict()
        d.update(overrides)
        cfg = ModelConfig(**d)
    return cfg
