"""Hugging Face Llama checkpoints <-> this framework's parameter names.

The reference has no model and therefore no weight format; users arriving with pretrained weights
have them in the ``transformers`` layout. The two layouts differ only in packing: this framework fuses
``q_proj | k_proj | v_proj`` into one ``wqkv`` (rows in that order) and ``gate_proj | up_proj`` into one
``w_gate_up``, pads the vocabulary to a multiple of 256, and uses the same rotate-half RoPE convention as
``transformers`` — so the conversion is concatenation, splitting and zero-padding; no permutation of rows.

    named = from_hf_state_dict(hf_model.state_dict(), cfg)
    engine.load_full_params(named)                       # every rank keeps its own rows

Checked against ``transformers.LlamaForCausalLM`` itself (tests/test_hf_compat.py): identical logits from the
``nn.Module`` model, identical loss from the native executor. One thing to know: the padded vocabulary rows are
zero, so each padded id contributes ``exp(0)`` to the softmax partition function until training pushes it down —
with 128 256 ids Llama-3 needs no padding at all, a 32 000-id vocabulary none either (both multiples of 256).
"""
from __future__ import annotations

from typing import Dict, Mapping

import torch

from .configs import ModelConfig


def config_from_hf(hf_cfg) -> ModelConfig:
    """``transformers.LlamaConfig`` -> ``ModelConfig``."""
    heads = hf_cfg.num_attention_heads
    rope = getattr(hf_cfg, "rope_theta", None)
    if rope is None:
        rope = (getattr(hf_cfg, "rope_parameters", None) or {}).get("rope_theta", 10000.0)
    return ModelConfig("llama", hf_cfg.vocab_size, hf_cfg.hidden_size, hf_cfg.num_hidden_layers, heads,
                       getattr(hf_cfg, "num_key_value_heads", None) or heads, hf_cfg.intermediate_size,
                       hf_cfg.max_position_embeddings, rope_theta=float(rope), norm_eps=float(hf_cfg.rms_norm_eps),
                       tie_embeddings=bool(getattr(hf_cfg, "tie_word_embeddings", False)))


def _pad_vocab(t: torch.Tensor, rows: int) -> torch.Tensor:
    if t.shape[0] == rows:
        return t
    out = torch.zeros((rows,) + tuple(t.shape[1:]), dtype=t.dtype)
    out[: t.shape[0]] = t
    return out


def from_hf_state_dict(sd: Mapping[str, torch.Tensor], cfg: ModelConfig) -> Dict[str, torch.Tensor]:
    g = lambda k: sd[k].detach().to(torch.float32).cpu()  # noqa: E731
    out: Dict[str, torch.Tensor] = {}
    out["embed.weight"] = _pad_vocab(g("model.embed_tokens.weight"), cfg.padded_vocab)
    out["final_norm.weight"] = g("model.norm.weight")
    head = "lm_head.weight" if "lm_head.weight" in sd else "model.embed_tokens.weight"      # tied embeddings
    out["lm_head.weight"] = _pad_vocab(g(head), cfg.padded_vocab)
    for i in range(cfg.n_layers):
        h, p = f"model.layers.{i}.", f"blocks.{i}."
        out[p + "attn_norm.weight"] = g(h + "input_layernorm.weight")
        out[p + "wqkv.weight"] = torch.cat([g(h + "self_attn.q_proj.weight"), g(h + "self_attn.k_proj.weight"),
                                            g(h + "self_attn.v_proj.weight")], 0)
        out[p + "wo.weight"] = g(h + "self_attn.o_proj.weight")
        out[p + "ffn_norm.weight"] = g(h + "post_attention_layernorm.weight")
        out[p + "w_gate_up.weight"] = torch.cat([g(h + "mlp.gate_proj.weight"), g(h + "mlp.up_proj.weight")], 0)
        out[p + "w_down.weight"] = g(h + "mlp.down_proj.weight")
    return out


def to_hf_state_dict(named: Mapping[str, torch.Tensor], cfg: ModelConfig) -> Dict[str, torch.Tensor]:
    """Inverse of :func:`from_hf_state_dict` (vocabulary padding rows are dropped)."""
    d, hd = cfg.hidden, cfg.head_dim
    nq, nk = cfg.n_heads * hd, cfg.n_kv_heads * hd
    out: Dict[str, torch.Tensor] = {}
    out["model.embed_tokens.weight"] = named["embed.weight"][: cfg.vocab_size].clone()
    out["model.norm.weight"] = named["final_norm.weight"].clone()
    if not cfg.tie_embeddings:
        out["lm_head.weight"] = named["lm_head.weight"][: cfg.vocab_size].clone()
    for i in range(cfg.n_layers):
        h, p = f"model.layers.{i}.", f"blocks.{i}."
        qkv = named[p + "wqkv.weight"]
        assert qkv.shape == (nq + 2 * nk, d)
        out[h + "self_attn.q_proj.weight"] = qkv[:nq].clone()
        out[h + "self_attn.k_proj.weight"] = qkv[nq:nq + nk].clone()
        out[h + "self_attn.v_proj.weight"] = qkv[nq + nk:].clone()
        out[h + "self_attn.o_proj.weight"] = named[p + "wo.weight"].clone()
        out[h + "input_layernorm.weight"] = named[p + "attn_norm.weight"].clone()
        out[h + "post_attention_layernorm.weight"] = named[p + "ffn_norm.weight"].clone()
        gu = named[p + "w_gate_up.weight"]
        out[h + "mlp.gate_proj.weight"] = gu[: cfg.ffn].clone()
        out[h + "mlp.up_proj.weight"] = gu[cfg.ffn:].clone()
        out[h + "mlp.down_proj.weight"] = named[p + "w_down.weight"].clone()
    return out


# ------------------------------------------------------------------ GPT-2
def gpt2_config_from_hf(hf_cfg) -> ModelConfig:
    """``transformers.GPT2Config`` -> ``ModelConfig`` (GELU must be the tanh form, "gelu_new")."""
    inner = hf_cfg.n_inner or 4 * hf_cfg.n_embd
    return ModelConfig("gpt2", hf_cfg.vocab_size, hf_cfg.n_embd, hf_cfg.n_layer, hf_cfg.n_head, hf_cfg.n_head, inner,
                       hf_cfg.n_positions, norm_eps=float(hf_cfg.layer_norm_epsilon), tie_embeddings=True)


_GPT2_LINEAR = (("attn.c_attn", "attn_qkv"), ("attn.c_proj", "attn_proj"), ("mlp.c_fc", "mlp_fc"), ("mlp.c_proj", "mlp_proj"))


def gpt2_from_hf_state_dict(sd: Mapping[str, torch.Tensor], cfg: ModelConfig) -> Dict[str, torch.Tensor]:
    """GPT-2 in ``transformers`` stores its projections as ``Conv1D`` ([in, out]); ours are ``nn.Linear``
    ([out, in]) — the conversion is a transpose. q | k | v are fused along the output dimension in both."""
    g = lambda k: sd[k].detach().to(torch.float32).cpu()  # noqa: E731
    out = {"wte.weight": _pad_vocab(g("transformer.wte.weight"), cfg.padded_vocab), "wpe.weight": g("transformer.wpe.weight"),
           "ln_f.weight": g("transformer.ln_f.weight"), "ln_f.bias": g("transformer.ln_f.bias")}
    for i in range(cfg.n_layers):
        h, p = f"transformer.h.{i}.", f"blocks.{i}."
        for ln in ("ln_1", "ln_2"):
            out[p + ln + ".weight"], out[p + ln + ".bias"] = g(h + ln + ".weight"), g(h + ln + ".bias")
        for theirs, ours in _GPT2_LINEAR:
            out[p + ours + ".weight"] = g(h + theirs + ".weight").t().contiguous()
            out[p + ours + ".bias"] = g(h + theirs + ".bias")
    return out
