"""GPT-2 as a plain ``torch.nn.Module`` (pre-LN, learned positions, GELU MLP, tied head).

Used by the generic ZeRO engine (BASELINE.json configs 1 and 2) on CPU/gloo and CUDA/NCCL.
The reference contains no model."""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import gpt2_ops as G2
from .configs import ModelConfig


class GPT2Block(nn.Module):
    def __init__(self, cfg: ModelConfig):
        super().__init__()
        d = cfg.hidden
        self.n_heads = cfg.n_heads
        self.ln_1 = nn.LayerNorm(d, eps=cfg.norm_eps)
        self.attn_qkv = nn.Linear(d, 3 * d)
        self.attn_proj = nn.Linear(d, d)
        self.ln_2 = nn.LayerNorm(d, eps=cfg.norm_eps)
        self.mlp_fc = nn.Linear(d, cfg.ffn)
        self.mlp_proj = nn.Linear(cfg.ffn, d)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        B, S, D = x.shape
        own = G2.enabled_for(x)       # LayerNorm / bias+GELU (csrc/gpt2_ops.cu) and the tcgen05 GEMM instead of torch ops
        lin = (lambda m, t: G2.linear(t, m.weight, m.bias)) if own else (lambda m, t: m(t))
        h = G2.layer_norm(x, self.ln_1.weight, self.ln_1.bias, self.ln_1.eps) if own else self.ln_1(x)
        q, k, v = lin(self.attn_qkv, h).view(B, S, 3, self.n_heads, D // self.n_heads).unbind(2)
        a = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=True)
        x = x + lin(self.attn_proj, a.transpose(1, 2).reshape(B, S, D))
        if own:
            h = G2.layer_norm(x, self.ln_2.weight, self.ln_2.bias, self.ln_2.eps)
            u = G2.bias_gelu(G2.linear(h, self.mlp_fc.weight), self.mlp_fc.bias)
        else:
            u = F.gelu(self.mlp_fc(self.ln_2(x)), approximate="tanh")
        return x + lin(self.mlp_proj, u)


class GPT2(nn.Module):
    def __init__(self, cfg: ModelConfig):
        super().__init__()
        self.cfg = cfg
        self.wte = nn.Embedding(cfg.padded_vocab, cfg.hidden)
        self.wpe = nn.Embedding(cfg.max_seq, cfg.hidden)
        self.blocks = nn.ModuleList(GPT2Block(cfg) for _ in range(cfg.n_layers))
        self.ln_f = nn.LayerNorm(cfg.hidden, eps=cfg.norm_eps)
        self.activation_checkpointing = False
        self.apply(self._init)
        for blk in self.blocks:  # GPT-2 residual-projection scaling
            for lin in (blk.attn_proj, blk.mlp_proj):
                nn.init.normal_(lin.weight, std=0.02 / math.sqrt(2 * cfg.n_layers))

    @staticmethod
    def _init(mod: nn.Module) -> None:
        if isinstance(mod, (nn.Linear, nn.Embedding)):
            nn.init.normal_(mod.weight, std=0.02)
        if isinstance(mod, nn.Linear) and mod.bias is not None:
            nn.init.zeros_(mod.bias)

    def forward(self, ids: torch.Tensor, labels: torch.Tensor = None):
        B, S = ids.shape
        x = self.wte(ids) + self.wpe(torch.arange(S, device=ids.device))[None]
        for blk in self.blocks:
            if self.activation_checkpointing and self.training:
                x = torch.utils.checkpoint.checkpoint(blk, x, use_reentrant=False)
            else:
                x = blk(x)
        logits = F.linear(self.ln_f(x), self.wte.weight)
        if labels is None:
            return logits
        return F.cross_entropy(logits.float().view(B * S, -1), labels.reshape(-1), ignore_index=-100)
