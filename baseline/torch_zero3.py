"""Library baseline for the headline benchmark (``bench.py --impl torch``).

The reference cannot be run on this box: its public API only writes a DeepSpeed JSON and ``Popen``s the
``deepspeed`` CLI (ref ``ai_engine/deepspeed_launcher.py:302-367``), DeepSpeed is not installable offline, and the
reference ships no training script. SURVEY.md §6 names the stand-in: what that DeepSpeed job would execute, written
with stock PyTorch only — cuBLAS GEMMs, cuDNN/flash SDPA attention, NCCL all-gather / reduce-scatter (FSDP2
``fully_shard`` = ZeRO-3: parameters, gradients and optimizer state sharded, each block gathered just in time in
forward and backward, prefetched one block ahead), fp32 master weights with bf16 compute and bf16 gradient
reduction, global-norm clipping at 1.0 and ``torch.optim.AdamW(fused=True)``.

Nothing from this repository's engine, models or kernels is on this path. The model below is a plain Llama-3
module written for speed in stock PyTorch (fused ``F.rms_norm``, SDPA with native GQA, cached RoPE tables) so the
comparison is against a well-written library program, not a straw man.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


class _Block(nn.Module):
    def __init__(self, c):
        super().__init__()
        d, hd = c.hidden, c.head_dim
        self.c = c
        self.attn_norm = nn.Parameter(torch.ones(d))
        self.wqkv = nn.Linear(d, (c.n_heads + 2 * c.n_kv_heads) * hd, bias=False)
        self.wo = nn.Linear(d, d, bias=False)
        self.ffn_norm = nn.Parameter(torch.ones(d))
        self.w_gate_up = nn.Linear(d, 2 * c.ffn, bias=False)
        self.w_down = nn.Linear(c.ffn, d, bias=False)

    def forward(self, x, cos, sin):
        c = self.c
        B, S, D = x.shape
        hd, Hq, Hk = c.head_dim, c.n_heads, c.n_kv_heads
        qkv = self.wqkv(F.rms_norm(x, (D,), self.attn_norm, c.norm_eps))
        q, k, v = qkv.split([Hq * hd, Hk * hd, Hk * hd], dim=-1)
        q = _rope(q.view(B, S, Hq, hd), cos, sin).transpose(1, 2)
        k = _rope(k.view(B, S, Hk, hd), cos, sin).transpose(1, 2)
        v = v.view(B, S, Hk, hd).transpose(1, 2)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=(Hq != Hk))
        x = x + self.wo(a.transpose(1, 2).reshape(B, S, D))
        g, u = self.w_gate_up(F.rms_norm(x, (D,), self.ffn_norm, c.norm_eps)).chunk(2, dim=-1)
        return x + self.w_down(F.silu(g) * u)


def _rope(x, cos, sin):
    a, b = x.chunk(2, dim=-1)
    return torch.cat([a * cos - b * sin, b * cos + a * sin], -1)


class TorchLlama(nn.Module):
    def __init__(self, c, seq_len: int):
        super().__init__()
        self.c = c
        self.embed = nn.Embedding(c.padded_vocab, c.hidden)
        self.blocks = nn.ModuleList(_Block(c) for _ in range(c.n_layers))
        self.final_norm = nn.Parameter(torch.ones(c.hidden))
        self.lm_head = nn.Linear(c.hidden, c.padded_vocab, bias=False)
        self.checkpoint_blocks = False
        hd = c.head_dim
        inv = 1.0 / (c.rope_theta ** (torch.arange(0, hd, 2, dtype=torch.float64) / hd))
        ang = torch.outer(torch.arange(seq_len, dtype=torch.float64), inv)
        self.register_buffer("cos", ang.cos().float()[None, :, None, :], persistent=False)
        self.register_buffer("sin", ang.sin().float()[None, :, None, :], persistent=False)

    def forward(self, ids, labels):
        x = self.embed(ids)
        cos, sin = self.cos.to(x.dtype), self.sin.to(x.dtype)
        for blk in self.blocks:
            if self.checkpoint_blocks:
                x = torch.utils.checkpoint.checkpoint(blk, x, cos, sin, use_reentrant=False)
            else:
                x = blk(x, cos, sin)
        logits = self.lm_head(F.rms_norm(x, (x.shape[-1],), self.final_norm, self.c.norm_eps))
        return F.cross_entropy(logits.view(-1, logits.shape[-1]).float(), labels.reshape(-1), ignore_index=-100)


def pick_checkpointing(c, world: int, micro_batch: int, seq_len: int, hbm_bytes: Optional[int]) -> bool:
    """Deterministic (same answer on every rank): per-block activation checkpointing — which the reference's config
    turns on by default (ref ``deepspeed_launcher.py:65,215-222``) — is used when the estimated footprint without
    it would not leave a 15 % margin of HBM."""
    if hbm_bytes is None:
        return False
    T = micro_batch * seq_len
    n = c.num_params()
    states = 16.0 * n / world                                   # fp32 master + fp32 grad + two fp32 moments, sharded
    qkv = (c.n_heads + 2 * c.n_kv_heads) * c.head_dim
    per_block = T * (8 * c.hidden + 2 * qkv + 4 * c.ffn) * 2.0   # what autograd keeps for one block (bf16)
    head = T * c.padded_vocab * (2 + 4 + 4)                     # bf16 logits, fp32 copy, fp32 gradient
    gathered = 3 * (n / max(c.n_layers, 1)) * 2.0               # blocks in flight, unsharded bf16
    return states + c.n_layers * per_block + head + gathered > 0.85 * hbm_bytes


class TorchZero3:
    """``train_step(ids, labels) -> {"loss": float}`` with the same semantics as the native engine's."""

    comm_mode = "nccl (FSDP2 fully_shard)"

    def __init__(self, mcfg, *, seq_len: int, micro_batch: int, device: torch.device, lr: float = 1e-4,
                 weight_decay: float = 0.01, grad_clip: float = 1.0, checkpoint_blocks: Optional[bool] = None,
                 seed: int = 1234):
        from torch.distributed.fsdp import MixedPrecisionPolicy, fully_shard
        self.device = device
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.grad_clip = grad_clip
        torch.manual_seed(seed)
        on_cuda = device.type == "cuda"
        with torch.device("meta"):
            model = TorchLlama(mcfg, seq_len)
        if checkpoint_blocks is None:
            hbm = torch.cuda.get_device_properties(device).total_memory if on_cuda else None
            checkpoint_blocks = pick_checkpointing(mcfg, self.world, micro_batch, seq_len, hbm)
        model.checkpoint_blocks = bool(checkpoint_blocks)
        self.recompute = "torch.utils.checkpoint per block" if checkpoint_blocks else "none"
        compute = torch.bfloat16 if on_cuda else torch.float32
        mp = MixedPrecisionPolicy(param_dtype=compute, reduce_dtype=compute)
        for blk in model.blocks:
            fully_shard(blk, mp_policy=mp)
        fully_shard(model, mp_policy=mp)
        model.to_empty(device=device)                           # sharded fp32 parameters, never the whole model
        with torch.no_grad():
            for name, p in model.named_parameters():
                if name.endswith("norm"):
                    p.fill_(1.0)
                else:
                    nn.init.normal_(p, std=0.02)
            hd = mcfg.head_dim
            inv = 1.0 / (mcfg.rope_theta ** (torch.arange(0, hd, 2, dtype=torch.float64, device=device) / hd))
            ang = torch.outer(torch.arange(seq_len, dtype=torch.float64, device=device), inv)
            model.cos = ang.cos().float()[None, :, None, :]
            model.sin = ang.sin().float()[None, :, None, :]
        self.model = model
        self.opt = torch.optim.AdamW(model.parameters(), lr=lr, betas=(0.9, 0.999), eps=1e-8,
                                     weight_decay=weight_decay, fused=on_cuda)
        self.global_steps = 0

    def train_step(self, ids: torch.Tensor, labels: torch.Tensor) -> Dict[str, float]:
        loss = self.model(ids, labels)
        loss.backward()
        norm = torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.grad_clip)
        self.opt.step()
        self.opt.zero_grad(set_to_none=True)
        self.global_steps += 1
        stats = torch.stack([loss.detach().float(), _local(norm).float()])
        if self.world > 1:
            mean = stats[0:1].clone()
            dist.all_reduce(mean)
            stats[0] = mean[0] / self.world
        loss_v, norm_v = stats.tolist()                         # the step's one host read (8 bytes)
        return {"loss": loss_v, "grad_norm": norm_v, "step": self.global_steps}


def _local(t):
    return t.full_tensor() if hasattr(t, "full_tensor") else t
