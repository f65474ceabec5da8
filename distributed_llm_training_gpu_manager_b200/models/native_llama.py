"""Llama executed directly on the sm_100a kernels: flat ZeRO buffers, hand-written forward and
backward (no autograd graph, no tracing compiler), per-layer activation recompute, weights
gathered per block (ZeRO-3) and gradients reduce-scattered per block as soon as they exist.

Per block (T = micro_batch·seq tokens)::

    h1  = rmsnorm(x)                    K5          qkv = h1 · Wqkvᵀ        K4 (tcgen05 GEMM)
    rope(q,k) in place                  K7          a   = flash_attn(qkv)   K6
    x1  = x + a · Woᵀ  ; h2 = rmsnorm(x1)   K4 + fused add-norm K5
    gu  = h2 · Wguᵀ ; act = silu(g)·u   K4 + K8     x2  = x1 + act · Wdownᵀ K4

Backward recomputes the block from its saved input, then runs dgrad (B operand MN-major) and
wgrad (both operands MN-major) GEMMs whose epilogues write straight into the block's slice of
the gradient bucket. The LM head never materialises full logits: token chunks go
GEMM → in-place softmax-CE (K9) → dgrad/wgrad while the chunk is still in L2.

The nn.Module twin in ``models/llama.py`` is the numerical oracle (same fused weight layout).
The reference ships no model.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Tuple, Union

import torch
import torch.distributed as dist

from ..engine.config import EngineConfig, load_config
from ..engine.zero import ParamSource, ZeroEngine
from ..ops import attention as A
from ..ops import elementwise as E
from ..ops import gemm as G
from .configs import ModelConfig


class LlamaParams(ParamSource):
    def __init__(self, cfg: ModelConfig, seed: int = 1234, init_device: Union[str, torch.device] = "cpu"):
        self.cfg, self.seed, self.init_device = cfg, seed, torch.device(init_device)
        self.unit_names = ["root"] + [f"block{i}" for i in range(cfg.n_layers)]

    def specs(self):
        c = self.cfg
        d, hd = c.hidden, c.head_dim
        qkv = (c.n_heads + 2 * c.n_kv_heads) * hd
        out = [("embed.weight", (c.padded_vocab, d), 0), ("final_norm.weight", (d,), 0),
               ("lm_head.weight", (c.padded_vocab, d), 0)]
        for i in range(c.n_layers):
            p = f"blocks.{i}."
            out += [(p + "attn_norm.weight", (d,), i + 1), (p + "wqkv.weight", (qkv, d), i + 1),
                    (p + "wo.weight", (d, d), i + 1), (p + "ffn_norm.weight", (d,), i + 1),
                    (p + "w_gate_up.weight", (2 * c.ffn, d), i + 1), (p + "w_down.weight", (d, c.ffn), i + 1)]
        return out

    def init_rows(self, name: str, shape, r0: int, r1: int) -> torch.Tensor:
        if name.endswith("norm.weight"):
            return torch.ones(shape, dtype=torch.float32)[r0:r1]
        # N(0, 0.02) per row block, seeded by (name, row-chunk) so any world size sees the same tensor
        cols = int(shape[1])
        out = torch.empty((r1 - r0, cols), dtype=torch.float32, device=self.init_device)
        CH = 1024
        base = (hash_name(name) ^ self.seed) & 0x7FFFFFFF
        r = r0
        while r < r1:
            c0 = r // CH
            lo, hi = c0 * CH, min((c0 + 1) * CH, int(shape[0]))
            g = torch.Generator(device=self.init_device).manual_seed(base * 4099 + c0)
            blk = torch.randn((hi - lo, cols), generator=g, device=self.init_device, dtype=torch.float32) * 0.02
            take_hi = min(hi, r1)
            out[r - r0: take_hi - r0] = blk[r - lo: take_hi - lo]
            r = take_hi
        return out


def activation_bytes(c: ModelConfig, micro_batch: int, seq_len: int, head_chunk: int = 2048) -> Dict[str, int]:
    """bf16 bytes of activations alive at the forward/backward turn-around, per recompute policy."""
    T = micro_batch * seq_len
    qkv = (c.n_heads + 2 * c.n_kv_heads) * c.head_dim
    ckpt = T * c.hidden * 2                                       # block input
    sel = T * (4 * c.hidden + qkv) * 2 + ckpt                     # + h1, attn, x1, h2, qkv
    gu = T * 2 * c.ffn * 2
    act = T * c.ffn * 2
    work = T * (2 * c.ffn * 3 + 6 * c.hidden + 2 * qkv) * 2 + head_chunk * c.padded_vocab * 2 * 2 + T * c.hidden * 4
    L = c.n_layers
    return {"full": L * ckpt + work + sel + gu + act, "selective": L * sel + work + gu + act,
            "light": L * (sel + gu) + work + act, "none": L * (sel + gu + act) + work}


def static_bytes(c: ModelConfig, world_size: int, stage: int = 3, offload_optimizer: bool = False,
                 grad_bytes: int = 2, offload_param: bool = False, param_slots: int = 2) -> Dict[str, float]:
    """Per-GPU bytes of the engine's persistent buffers (ZeRO-3 native path), for capacity planning
    against 180 GB of HBM3e (BASELINE.json config 5: Llama-3-70B on 8 × B200)."""
    d, hd = c.hidden, c.head_dim
    qkv = (c.n_heads + 2 * c.n_kv_heads) * hd
    block = d * (qkv + d) + 3 * d * c.ffn + 2 * d
    root = 2 * c.padded_vocab * d + d
    total = c.n_layers * block + root
    shard = total / max(world_size, 1) if stage >= 1 else total
    # offload_param: the 16-bit shard lives in pinned host memory; the device keeps one staging buffer per resident
    # block plus one for the root unit (a unit's rows of the shard)
    staging = 2.0 * (param_slots * block + root) / max(world_size, 1)
    out = {"optimizer_states": 0.0 if offload_optimizer else 12.0 * shard,
           "param_shard": staging if (offload_param and stage >= 3) else 2.0 * shard,
           "grad_shard": float(grad_bytes) * shard,
           "param_buffers": 2.0 * (max(param_slots, 2) * block + root) if stage >= 3 else 2.0 * total,
           "grad_buffers": 2.0 * (2 * block + root) if stage >= 2 else 2.0 * total,
           "host_pinned": (12.0 * shard if offload_optimizer else 0.0) + (2.0 * shard if offload_param else 0.0),
           "params_total": float(total)}
    out["device_total"] = sum(v for k, v in out.items() if k not in ("host_pinned", "params_total"))
    return out


def hash_name(name: str) -> int:
    h = 2166136261
    for ch in name.encode():
        h = ((h ^ ch) * 16777619) & 0xFFFFFFFF
    return h


class ActivationOffloader:
    """``activation_checkpointing.cpu_checkpointing`` (ref ``deepspeed_launcher.py:67,218``; on in the
    reference's 70b preset): the tensors a block keeps for backward are copied to pinned host memory
    on a side stream as soon as the block has produced them, and brought back one block ahead of the
    backward pass, so HBM holds at most two blocks' worth of saved activations."""

    def __init__(self, device: torch.device):
        self.dev = device
        self.stream = torch.cuda.Stream(device=device)
        self.host: Dict[int, Dict[str, torch.Tensor]] = {}
        self.meta: Dict[int, Dict[str, Any]] = {}
        self.back: Dict[int, Tuple[Dict[str, Any], Any]] = {}
        self._pool: Dict[Tuple, List[torch.Tensor]] = {}

    def _pinned(self, like: torch.Tensor) -> torch.Tensor:
        key = (tuple(like.shape), like.dtype)
        free = self._pool.setdefault(key, [])
        return free.pop() if free else torch.empty(like.shape, dtype=like.dtype, pin_memory=True)

    def stash(self, l: int, tensors: Dict[str, Any]) -> None:
        """Start the device->host copy of every tensor in ``tensors`` (non-tensors are kept as is)."""
        self.stream.wait_stream(torch.cuda.current_stream())
        host, meta = {}, {}
        with torch.cuda.stream(self.stream):
            for k, v in tensors.items():
                if torch.is_tensor(v) and v.is_cuda:
                    buf = self._pinned(v)
                    buf.copy_(v, non_blocking=True)
                    v.record_stream(self.stream)          # keep the device block alive until the copy ran
                    host[k] = buf
                else:
                    meta[k] = v
        self.host[l], self.meta[l] = host, meta

    def prefetch(self, l: int) -> None:
        if l in self.back or l not in self.host:
            return
        out = dict(self.meta[l])
        with torch.cuda.stream(self.stream):
            for k, buf in self.host[l].items():
                out[k] = buf.to(self.dev, non_blocking=True)
            evt = torch.cuda.Event()
            evt.record(self.stream)
        self.back[l] = (out, evt)

    def fetch(self, l: int) -> Dict[str, Any]:
        self.prefetch(l)
        out, evt = self.back.pop(l)
        torch.cuda.current_stream().wait_event(evt)
        for t in out.values():
            if torch.is_tensor(t) and t.is_cuda:
                t.record_stream(torch.cuda.current_stream())
        for k, buf in self.host.pop(l).items():
            self._pool.setdefault((tuple(buf.shape), buf.dtype), []).append(buf)
        self.meta.pop(l, None)
        return out


class NativeLlamaEngine(ZeroEngine):
    """ZeRO engine whose compute path is the hand-written kernel sequence above."""

    def __init__(self, model_cfg: ModelConfig, config: Union[str, Dict[str, Any], EngineConfig], *, seq_len: int,
                 device: Optional[torch.device] = None, comm: str = "auto", seed: int = 1234,
                 attn_backend: Optional[str] = None, head_chunk: int = 2048, recompute: Optional[str] = None):
        assert model_cfg.family == "llama"
        ecfg = load_config(config)
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        self.mcfg = model_cfg
        self.seq_len = seq_len
        self.attn_backend = attn_backend
        self.head_chunk = head_chunk
        src = LlamaParams(model_cfg, seed, init_device=device if torch.device(device).type == "cuda" else "cpu")
        super().__init__(src, ecfg, device=device, comm=comm, broadcast_init=False)
        assert self.compute_dtype == torch.bfloat16 or not self.on_cuda, "native kernels are bf16"
        self._saved: Optional[Dict[str, Any]] = None
        self._micro_in_window = 0
        # Activation policy (``activation_checkpointing`` block of the config):
        #   full      — keep only each block's input, recompute the whole block in backward
        #   selective — keep everything except the two largest tensors (gate|up and silu·up, 0.7 GB per
        #               block at 8k tokens) and recompute just that GEMM + SwiGLU; default when the
        #               config enables activation checkpointing (180 GB HBM3e leaves room for the rest)
        #   light     — keep everything except silu·up; backward re-runs only the SwiGLU kernel
        #   none      — keep everything (config has no activation_checkpointing block)
        #   auto      — (default with activation checkpointing) the cheapest policy whose saved activations
        #               fit in the HBM left after building the engine with a 12 GB margin
        import os as _os
        self.recompute = recompute or _os.environ.get("B200ZERO_RECOMPUTE") or ("auto" if self.cfg.act_ckpt else "none")
        assert self.recompute in ("full", "selective", "light", "none", "auto")
        if self.recompute == "auto":
            self.recompute = self._pick_recompute_policy()
        self._zero_after_reduce = False     # every gradient is overwritten by its producer (see _reduce_unit)
        # K2: with the fused backend at ZeRO-3 every forward GEMM gathers its own weight in-kernel
        import os
        self._ag_fused = (self.fused is not None and self.stage >= 3 and self.pooled_params
                          and os.environ.get("B200ZERO_AG_GEMM", "1") == "1")
        self._ag_ptrs: Dict[str, Any] = {}
        # K2 dgrad flavour (opt-in: validated, but 1-3 % slower end to end): with nothing recomputed every weight is used
        # exactly once in backward — by its dgrad — so that GEMM gathers it in-kernel and backward needs no
        # bulk gather / prefetch stream at all.
        self._ag_dgrad = (self._ag_fused and self.recompute in ("none", "light")
                          and os.environ.get("B200ZERO_AG_DGRAD", "0") == "1")
        self._act_offload = (ActivationOffloader(self.device)
                             if (self.cfg.cpu_checkpointing and self.on_cuda) else None)

    def _pick_recompute_policy(self) -> str:
        """Cheapest-recompute policy whose saved activations fit in the HBM left after the engine's
        static buffers, with a 12 GB margin: none → light → selective → full."""
        if not self.on_cuda:
            return "selective"
        need = activation_bytes(self.mcfg, self.cfg.micro_batch, self.seq_len, self.head_chunk)
        free = torch.cuda.mem_get_info(self.device)[0] + (torch.cuda.memory_reserved(self.device)
                                                         - torch.cuda.memory_allocated(self.device))
        order = ["none", "light", "selective", "full"]
        pick = next((i for i, p in enumerate(order) if free - need[p] > 12 * 2 ** 30), len(order) - 1)
        if self.world_size > 1:                                  # all ranks must run the same kernel sequence
            t = torch.tensor([pick], device=self.device, dtype=torch.int32)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            pick = int(t.item())
        return order[pick]

    # -------------------------------------------------------------- helpers
    def _w(self, layer: int, name: str) -> torch.Tensor:
        return self.param_view(f"blocks.{layer}.{name}.weight")

    def _gw(self, layer: int, name: str) -> torch.Tensor:
        return self.grad_view(f"blocks.{layer}.{name}.weight")

    def _unit_is_resident(self, ui: int) -> bool:
        return self._resident[ui] and (ui == 0 or self._slot_owner[self._pslot(ui)] == ui)

    def _linear(self, l: int, name: str, x: torch.Tensor, ag: bool) -> torch.Tensor:
        """y = x · Wᵀ; when ``ag`` the weight is all-gathered inside the GEMM kernel (K2) and left in
        the block's gathered buffer for the rest of forward / recompute / backward."""
        w = self._w(l, name)
        if not ag:
            return G.linear_fwd(x, w)
        from ..ops.ag_gemm import ag_linear_fwd
        full = f"blocks.{l}.{name}.weight"
        ptrs = self._ag_ptrs.get(full)
        if ptrs is None:
            spec = self.layout.by_name[full]
            assert spec.partitioned, f"{full} is not partitioned; cannot fuse its gather"
            ptrs = self.fused._ptrs_for(self.shard_params[spec.shard_offset:spec.shard_offset + spec.shard_numel])
            self._ag_ptrs[full] = ptrs
        return ag_linear_fwd(x, ptrs, w, self.rank, self.world_size)

    def _dgrad(self, l: int, name: str, dy: torch.Tensor, ag: bool) -> torch.Tensor:
        """dx = dy · W; when ``ag`` the weight is all-gathered inside the GEMM kernel (K2, dgrad flavour)."""
        w = self._w(l, name)
        if not ag:
            return G.linear_dgrad(dy, w)
        from ..ops.ag_gemm import ag_linear_dgrad
        full = f"blocks.{l}.{name}.weight"
        ptrs = self._ag_ptrs.get(full)
        if ptrs is None:
            spec = self.layout.by_name[full]
            ptrs = self.fused._ptrs_for(self.shard_params[spec.shard_offset:spec.shard_offset + spec.shard_numel])
            self._ag_ptrs[full] = ptrs
        return ag_linear_dgrad(dy, ptrs, w, self.rank, self.world_size)

    def _begin_block(self, l: int) -> bool:
        """Make block ``l``'s parameters available. Returns True when the GEMMs must gather their own
        weights (fused all-gather→GEMM); otherwise the unit is (made) resident by a bulk gather."""
        ui = l + 1
        if not self.pooled_params or self._unit_is_resident(ui):
            evt = self._prefetch_evt.pop(ui, None)
            if evt is not None:
                torch.cuda.current_stream().wait_event(evt)
            return False
        if not self._ag_fused:
            self._ensure_resident(ui)
            return False
        u = self.layout.units[ui]
        pb = self._unit_pbuf[ui]
        for spec in u.params:                      # replicated (persistent) tensors: local copies
            if not spec.partitioned:
                pb[spec.full_offset:spec.full_offset + spec.numel].copy_(
                    self.shard_params[spec.shard_offset:spec.shard_offset + spec.shard_numel])
        slot = self._pslot(ui)
        prev = self._slot_owner[slot]
        if prev is not None and prev != ui:
            self._resident[prev] = False
        self._slot_owner[slot] = ui
        self._resident[ui] = True                  # true once the four AG-GEMMs below have run (stream order)
        return True

    def _begin_block_bwd(self, l: int) -> None:
        """Backward needs every weight of the block resident (bulk gather, prefetched one block ahead)."""
        ui = l + 1
        if self.pooled_params:
            self._ensure_resident(ui)

    def _block_fwd(self, l: int, x: torch.Tensor, keep: bool) -> Tuple[torch.Tensor, Optional[Dict[str, Any]]]:
        c = self.mcfg
        B, S = self._B, self.seq_len
        Hq, Hk, hd, F_ = c.n_heads, c.n_kv_heads, c.head_dim, c.ffn
        ag = self._begin_block(l)
        h1, rstd1, _ = E.rmsnorm_fwd(x, self._w(l, "attn_norm"), c.norm_eps)
        qkv = self._linear(l, "wqkv", h1, ag)
        E.rope_(qkv, S, Hq + Hk, hd, c.rope_theta)
        attn, actx = A.attention_fwd(qkv, B, S, Hq, Hk, hd, backend=self.attn_backend)
        o = self._linear(l, "wo", attn, ag)
        h2, rstd2, x1 = E.rmsnorm_fwd(x, self._w(l, "ffn_norm"), c.norm_eps, delta=o)
        gu = self._linear(l, "w_gate_up", h2, ag)
        act = E.swiglu_fwd(gu)
        dn = self._linear(l, "w_down", act, ag)
        x2 = E.add(x1, dn, out=dn)
        if not keep:
            return x2, None
        return x2, dict(h1=h1, rstd1=rstd1, attn=attn, actx=actx, x1=x1, rstd2=rstd2, h2=h2, gu=gu, act=act)

    def _block_bwd(self, l: int, x: torch.Tensor, dx2: torch.Tensor, acc: bool,
                   sv: Optional[Dict[str, Any]] = None) -> torch.Tensor:
        c = self.mcfg
        S = self.seq_len
        Hq, Hk, hd = c.n_heads, c.n_kv_heads, c.head_dim
        agd = False
        if sv is None:
            _, sv = self._block_fwd(l, x, keep=True)                  # full recompute
        else:
            if self._ag_dgrad and "gu" in sv:
                agd = self._begin_block(l)                             # True: the dgrad GEMMs gather their own weights
            else:
                self._begin_block_bwd(l)
            if "gu" not in sv:                                        # selective: redo gate|up GEMM + SwiGLU only
                sv["gu"] = G.linear_fwd(sv["h2"], self._w(l, "w_gate_up"))
            if "act" not in sv:                                       # light: SwiGLU only
                sv["act"] = E.swiglu_fwd(sv["gu"])
        adt = dict(accumulate=acc)
        # MLP
        G.linear_wgrad(dx2, sv["act"], out=self._gw(l, "w_down"), **adt)
        dact = self._dgrad(l, "w_down", dx2, agd)
        dgu = E.swiglu_bwd(sv["gu"], dact)
        del dact
        G.linear_wgrad(dgu, sv["h2"], out=self._gw(l, "w_gate_up"), **adt)
        dh2 = self._dgrad(l, "w_gate_up", dgu, agd)
        del dgu
        dx1, _ = E.rmsnorm_bwd(sv["x1"], self._w(l, "ffn_norm"), dh2, c.norm_eps, rstd=sv["rstd2"], dres=dx2,
                               dw_out=self._gw(l, "ffn_norm"), dw_accumulate=acc, dx_out=dh2)
        # attention
        G.linear_wgrad(dx1, sv["attn"], out=self._gw(l, "wo"), **adt)
        dattn = self._dgrad(l, "wo", dx1, agd)
        dqkv = A.attention_bwd(sv["actx"], dattn)
        E.rope_(dqkv, S, Hq + Hk, hd, c.rope_theta, inverse=True)
        G.linear_wgrad(dqkv, sv["h1"], out=self._gw(l, "wqkv"), **adt)
        dh1 = self._dgrad(l, "wqkv", dqkv, agd)
        dx, _ = E.rmsnorm_bwd(x, self._w(l, "attn_norm"), dh1, c.norm_eps, rstd=sv["rstd1"], dres=dx1,
                              dw_out=self._gw(l, "attn_norm"), dw_accumulate=acc, dx_out=dh1)
        return dx

    @torch.no_grad()
    def eval_loss(self, ids: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """Mean token loss of one batch without touching gradients or saved activations (validation loss for the
        loss monitor). Same kernels as ``forward``; the LM head runs chunk by chunk without its dgrad / wgrad."""
        c = self.mcfg
        B, S = ids.shape
        assert S == self.seq_len, f"engine built for seq_len={self.seq_len}, got {S}"
        self._B = B
        T = B * S
        if self.pooled_params:
            self._ensure_resident(0)
        ids_flat, labels_flat = ids.reshape(-1), labels.reshape(-1)
        x = E.embedding_fwd(ids_flat, self.param_view("embed.weight"))
        for l in range(c.n_layers):
            if self.pooled_params and not self._ag_fused and l + 2 <= c.n_layers:
                self._prefetch(l + 2)
            x, _ = self._block_fwd(l, x, keep=False)
        hf, _, _ = E.rmsnorm_fwd(x, self.param_view("final_norm.weight"), c.norm_eps)
        w_head = self.param_view("lm_head.weight")
        loss_tok = torch.empty(T, dtype=torch.float32, device=hf.device)
        for t0 in range(0, T, self.head_chunk):
            t1 = min(T, t0 + self.head_chunk)
            logits = G.linear_fwd(hf[t0:t1], w_head)
            loss_tok[t0:t1] = E.cross_entropy_(logits, labels_flat[t0:t1], write_grad=False)
            del logits
        n_valid = (labels_flat >= 0).sum().clamp(min=1).float()
        return loss_tok.sum() / n_valid

    # -------------------------------------------------------------- protocol
    def forward(self, ids: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        c, cfg = self.mcfg, self.cfg
        self._in_backward = False
        B, S = ids.shape
        assert S == self.seq_len, f"engine built for seq_len={self.seq_len}, got {S}"
        self._B = B
        T = B * S
        L = c.n_layers
        acc = self.stage <= 1 and self._micro_in_window > 0
        with self.timers.phase("forward"):
            if self.pooled_params:
                self._ensure_resident(0)
            ids_flat, labels_flat = ids.reshape(-1), labels.reshape(-1)
            x = E.embedding_fwd(ids_flat, self.param_view("embed.weight"))
            ckpts: List[torch.Tensor] = []
            saved_acts: List[Optional[Dict[str, Any]]] = []
            for l in range(L):
                if self.pooled_params and not self._ag_fused and l + 2 <= L:
                    self._prefetch(l + 2)          # bulk gather of the next block on the side stream
                ckpts.append(x)
                x, sv_l = self._block_fwd(l, x, keep=self.recompute != "full")
                if sv_l is not None and self.recompute == "selective":
                    sv_l.pop("gu"); sv_l.pop("act")
                elif sv_l is not None and self.recompute == "light":
                    sv_l.pop("act")
                if self._act_offload is not None:
                    pack = dict(sv_l or {}, _x=ckpts[-1])
                    actx = pack.pop("actx", None)
                    if actx is not None and actx[0] == "native":      # (tag, qkv, out, lse, dims)
                        pack.update(_qkv=actx[1], _lse=actx[3], _dims=actx[4])
                    elif actx is not None:
                        pack["actx"] = actx                            # sdpa autograd context stays on device
                    self._act_offload.stash(l, pack)
                    ckpts[-1] = None
                    sv_l = {"_offloaded": True} if sv_l is not None else None
                saved_acts.append(sv_l)
            hf, rstd_f, _ = E.rmsnorm_fwd(x, self.param_view("final_norm.weight"), c.norm_eps)
            # LM head + CE, chunked over tokens; gradients of the head are produced here
            w_head, g_head = self.param_view("lm_head.weight"), self.grad_view("lm_head.weight")
            n_valid = (labels_flat >= 0).sum().clamp(min=1).float()
            dhf = torch.empty_like(hf)
            loss_tok = torch.empty(T, dtype=torch.float32, device=hf.device)
            gscale = self.scaler.scale / cfg.grad_accum
            inv_valid = float(1.0 / n_valid.item()) if not self.on_cuda else None
            if inv_valid is None:
                # every token is a target in the packed synthetic/pretraining setting; avoid a host sync
                inv_valid = 1.0 / T if self._all_valid else float(1.0 / n_valid.item())
            for ci, t0 in enumerate(range(0, T, self.head_chunk)):
                t1 = min(T, t0 + self.head_chunk)
                logits = G.linear_fwd(hf[t0:t1], w_head)
                loss_tok[t0:t1] = E.cross_entropy_(logits, labels_flat[t0:t1], grad_scale=gscale * inv_valid)
                G.linear_dgrad(logits, w_head, out=dhf[t0:t1])
                G.linear_wgrad(logits, hf[t0:t1], out=g_head, accumulate=(acc or ci > 0))
                del logits
            loss = loss_tok.sum() * inv_valid
        self._saved = dict(ids=ids_flat, ckpts=ckpts, acts=saved_acts, x_last=x, rstd_f=rstd_f, dhf=dhf, acc=acc)
        return loss

    _all_valid = True

    def backward(self, loss: torch.Tensor) -> None:
        cfg, c = self.cfg, self.mcfg
        sv = self._saved
        assert sv is not None, "backward() without a preceding forward()"
        self._in_backward = True
        boundary = self.is_gradient_accumulation_boundary()
        self._reduce_now = self.stage >= 2 or boundary
        self._reduced = [False] * len(self.layout.units)
        if self.fused is not None:
            self.fused.begin_micro_step()
        hook = self.fault_hooks.get("pre_backward")
        if hook is not None:
            loss = hook(self, loss)
        self._loss_accum += loss.detach().float() / cfg.grad_accum
        acc = sv["acc"]
        L = c.n_layers
        with self.timers.phase("backward"):
            dx, _ = E.rmsnorm_bwd(sv["x_last"], self.param_view("final_norm.weight"), sv["dhf"], c.norm_eps,
                                  rstd=sv["rstd_f"], dw_out=self.grad_view("final_norm.weight"), dw_accumulate=acc)
            for l in range(L - 1, -1, -1):
                ui = l + 1
                if (self.pooled_params and ui - 1 >= 1 and (not self._ag_fused or self.recompute != "full")
                        and not self._ag_dgrad):
                    self._prefetch(ui - 1)
                if self.pooled_grads and self.on_cuda:
                    evt = self._gslot_free_evt[(ui - 1) % 2]
                    if evt is not None:
                        torch.cuda.current_stream().wait_event(evt)
                x_l, acts_l = sv["ckpts"][l], sv["acts"][l]
                if self._act_offload is not None:
                    pack = self._act_offload.fetch(l)
                    if l - 1 >= 0:
                        self._act_offload.prefetch(l - 1)
                    x_l = pack.pop("_x")
                    if acts_l is not None:
                        if "_qkv" in pack:
                            pack["actx"] = ("native", pack.pop("_qkv"), pack["attn"], pack.pop("_lse"), pack.pop("_dims"))
                        acts_l = pack
                dx = self._block_bwd(l, x_l, dx, acc, acts_l)
                sv["ckpts"][l] = None
                sv["acts"][l] = None
                if self._reduce_now:
                    self._reduce_unit(ui)
            if not acc:
                self.grad_view("embed.weight").zero_()               # scatter-add target; everything else is overwritten
            E.embedding_bwd(sv["ids"], dx, self.grad_view("embed.weight"))
            if self._reduce_now:
                self._reduce_unit(0)
        if self.on_cuda and cfg.overlap_comm:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        hook = self.fault_hooks.get("post_backward")
        if hook is not None:
            hook(self)
        self._saved = None
        self._micro_in_window = 0 if boundary else self._micro_in_window + 1
        self._in_backward = False

    def train_step(self, ids: torch.Tensor, labels: torch.Tensor) -> Optional[Dict[str, float]]:
        loss = self.forward(ids, labels)
        self.backward(loss)
        return self.step()
