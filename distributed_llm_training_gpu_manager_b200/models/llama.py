"""Llama-3 as a plain ``torch.nn.Module`` — the fp32/bf16 *oracle* for the native sm_100a
executor (``models/native_llama.py``) and the body of the NCCL+cuBLAS baseline arm. Weights
use the same fused layouts as the native path (``wqkv`` = [q;k;v], ``w_gate_up`` = [gate;up])
so state dicts are interchangeable."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .configs import ModelConfig


def _rope(x: torch.Tensor, theta: float) -> torch.Tensor:
    """x: [B, S, H, hd], rotate-half convention."""
    B, S, H, hd = x.shape
    inv = 1.0 / (theta ** (torch.arange(0, hd, 2, dtype=torch.float64, device=x.device) / hd))
    ang = torch.outer(torch.arange(S, dtype=torch.float64, device=x.device), inv)
    c, s = ang.cos().float()[None, :, None, :], ang.sin().float()[None, :, None, :]
    a, b = x.float()[..., : hd // 2], x.float()[..., hd // 2:]
    return torch.cat([a * c - b * s, b * c + a * s], -1).to(x.dtype)


class RMSNorm(nn.Module):
    def __init__(self, d: int, eps: float):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(d))
        self.eps = eps

    def forward(self, x):
        xf = x.float()
        return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + self.eps) * self.weight.float()).to(x.dtype)


class LlamaBlock(nn.Module):
    def __init__(self, cfg: ModelConfig):
        super().__init__()
        d, hd = cfg.hidden, cfg.head_dim
        self.cfg = cfg
        self.attn_norm = RMSNorm(d, cfg.norm_eps)
        self.wqkv = nn.Linear(d, (cfg.n_heads + 2 * cfg.n_kv_heads) * hd, bias=False)
        self.wo = nn.Linear(d, d, bias=False)
        self.ffn_norm = RMSNorm(d, cfg.norm_eps)
        self.w_gate_up = nn.Linear(d, 2 * cfg.ffn, bias=False)
        self.w_down = nn.Linear(cfg.ffn, d, bias=False)

    def forward(self, x):
        c = self.cfg
        B, S, D = x.shape
        hd, Hq, Hk = c.head_dim, c.n_heads, c.n_kv_heads
        qkv = self.wqkv(self.attn_norm(x))
        q = _rope(qkv[..., : Hq * hd].reshape(B, S, Hq, hd), c.rope_theta)
        k = _rope(qkv[..., Hq * hd: (Hq + Hk) * hd].reshape(B, S, Hk, hd), c.rope_theta)
        v = qkv[..., (Hq + Hk) * hd:].reshape(B, S, Hk, hd)
        rep = Hq // Hk
        k, v = k.repeat_interleave(rep, 2), v.repeat_interleave(rep, 2)
        a = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=True)
        x = x + self.wo(a.transpose(1, 2).reshape(B, S, D))
        gu = self.w_gate_up(self.ffn_norm(x))
        return x + self.w_down(F.silu(gu[..., : c.ffn]) * gu[..., c.ffn:])


class Llama(nn.Module):
    def __init__(self, cfg: ModelConfig):
        super().__init__()
        self.cfg = cfg
        self.embed = nn.Embedding(cfg.padded_vocab, cfg.hidden)
        self.blocks = nn.ModuleList(LlamaBlock(cfg) for _ in range(cfg.n_layers))
        self.final_norm = RMSNorm(cfg.hidden, cfg.norm_eps)
        self.lm_head = nn.Linear(cfg.hidden, cfg.padded_vocab, bias=False)
        self.activation_checkpointing = False
        for m in self.modules():
            if isinstance(m, (nn.Linear, nn.Embedding)):
                nn.init.normal_(m.weight, std=0.02)

    def forward(self, ids, labels=None):
        B, S = ids.shape
        x = self.embed(ids)
        for blk in self.blocks:
            if self.activation_checkpointing and self.training:
                x = torch.utils.checkpoint.checkpoint(blk, x, use_reentrant=False)
            else:
                x = blk(x)
        logits = self.lm_head(self.final_norm(x))
        if labels is None:
            return logits
        return F.cross_entropy(logits.float().view(B * S, -1), labels.reshape(-1), ignore_index=-100)
