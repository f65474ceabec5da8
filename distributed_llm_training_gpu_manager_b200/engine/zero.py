"""ZeRO-0/1/2/3 training engine for arbitrary ``torch.nn.Module`` models.

This is the engine the reference only *configures*: its launcher emits a DeepSpeed JSON and
shells out (ref ``ai_engine/deepspeed_launcher.py:302-367``); nothing in that tree partitions,
reduces or steps anything. The user-facing protocol mirrors DeepSpeed's so a
training script ports by changing one import::

    engine = initialize(model=model, config="ds_config.json")
    loss = engine(ids, labels); engine.backward(loss); engine.step()
    engine.save_checkpoint(dir); engine.load_checkpoint(dir)

Design (see ``partition.py`` for the layout):

* One process per GPU, ``torch.distributed`` (NCCL on GPUs, gloo on CPU — the CPU tier runs the
  same code and is the oracle for the fused kernels).
* Parameters live in flat buffers; ``param.data`` and ``param.grad`` are *views* into them, bound
  once at construction. Stage ≤ 2 keeps every unit gathered; stage 3 keeps only the root unit and
  two block-sized buffers (even/odd blocks) alive, re-filled by all-gather just before a block
  runs (forward, recompute and backward) and prefetched one block ahead on a side stream.
* Gradients: autograd accumulates straight into the unit's gradient buffer; when the last
  parameter of a unit has its gradient (post-accumulate hooks), the unit is reduce-scattered
  into the rank's gradient shard on the communication stream (stage ≥ 2: every micro-step;
  stage ≤ 1: at the accumulation boundary).
* ``step()``: ‖g‖² + isfinite on the shard → one 3-float all-reduce (Σ‖g‖², non-finite flag,
  loss) → clip coefficient on device → fused AdamW predicated on the flag → bf16 write-back →
  (stage ≤ 2) all-gather of updated rows. The only host sync is the read of those 3 floats,
  which feeds :class:`LossSpikeMonitor` for the halt / rollback decision.
* ``comm="fused"`` swaps the NCCL reduce-scatter / all-gather for the in-kernel peer-memory
  versions (``parallel/fused.py``: K1/K3).
"""
from __future__ import annotations

import logging
import os
import time
from contextlib import contextmanager, nullcontext
from typing import Any, Callable, Dict, List, Optional, Tuple, Union

import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import optim as O
from .config import EngineConfig, load_config
from .partition import Layout, ParamSpec, build_layout
from .schedules import DynamicLossScaler, LRSchedule

_BLOCK_LIST_NAMES = ("blocks", "layers", "h", "transformer_blocks")
_log = logging.getLogger(__name__)


def _dist_ready() -> bool:
    return dist.is_available() and dist.is_initialized()


def discover_units(module: nn.Module) -> Tuple[List[str], List[nn.Module], Dict[int, int]]:
    """Unit 0 = root (everything outside the block list); units 1.. = transformer blocks."""
    blocks: List[nn.Module] = []
    for name in _BLOCK_LIST_NAMES:
        child = getattr(module, name, None)
        if isinstance(child, nn.ModuleList) and len(child) > 0:
            blocks = list(child)
            break
    unit_of: Dict[int, int] = {}
    for bi, blk in enumerate(blocks):
        for p in blk.parameters():
            unit_of.setdefault(id(p), bi + 1)
    names = ["root"] + [f"block{i}" for i in range(len(blocks))]
    return names, blocks, unit_of


class PhaseTimers:
    """CUDA-event (or perf_counter) timers per phase; ``wall_clock_breakdown`` keeps its meaning
    (ref ``deepspeed_launcher.py:79,129``)."""

    def __init__(self, enabled: bool, cuda: bool):
        self.enabled, self.cuda = enabled, cuda
        self.totals: Dict[str, float] = {}
        self._open: List[Tuple[str, Any, Any]] = []
        # B200ZERO_NVTX=1: forward / backward / reduce / allgather / step show up as NVTX ranges in nsys / ncu
        self.nvtx = bool(cuda) and os.environ.get("B200ZERO_NVTX", "0") == "1"

    @contextmanager
    def phase(self, name: str):
        if self.nvtx:
            torch.cuda.nvtx.range_push(name)
            try:
                with self._timed(name):
                    yield
            finally:
                torch.cuda.nvtx.range_pop()
        else:
            with self._timed(name):
                yield

    @contextmanager
    def _timed(self, name: str):
        if not self.enabled:
            yield
            return
        if self.cuda:
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            yield
            e.record()
            self._open.append((name, s, e))
        else:
            t0 = time.perf_counter()
            yield
            self.totals[name] = self.totals.get(name, 0.0) + (time.perf_counter() - t0) * 1e3

    def collect(self) -> Dict[str, float]:
        if self.cuda and self._open:
            torch.cuda.synchronize()
            for name, s, e in self._open:
                self.totals[name] = self.totals.get(name, 0.0) + s.elapsed_time(e)
            self._open.clear()
        out, self.totals = self.totals, {}
        return out


class ParamSource:
    """Parameter provider for engines that own their compute path (no ``nn.Module``): names,
    shapes and units, plus a deterministic row-range initialiser so a rank can fill its shard of
    a 70B model without ever materialising the whole tensor."""

    unit_names: List[str] = []

    def specs(self) -> List[Tuple[str, Tuple[int, ...], int]]:
        raise NotImplementedError

    def init_rows(self, name: str, shape: Tuple[int, ...], r0: int, r1: int) -> torch.Tensor:
        """fp32 tensor for rows [r0, r1) of parameter ``name`` (any device)."""
        raise NotImplementedError


class ZeroEngine:
    def __init__(self, module: nn.Module, config: Union[str, Dict[str, Any], EngineConfig], *,
                 device: Optional[torch.device] = None, comm: str = "auto", broadcast_init: bool = True):
        self.cfg = load_config(config)
        self.module = module
        self.rank = dist.get_rank() if _dist_ready() else 0
        self.world_size = dist.get_world_size() if _dist_ready() else 1
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        self.device = torch.device(device)
        self.on_cuda = self.device.type == "cuda"
        cfg = self.cfg
        self.stage = cfg.stage
        self.compute_dtype = cfg.compute_dtype
        self.global_steps = 0
        self.global_samples = 0
        self.micro_steps = 0
        self.skipped_steps = 0
        self.scaler = DynamicLossScaler(cfg.precision == "fp16", cfg.loss_scale, cfg.initial_scale_power,
                                        cfg.loss_scale_window, cfg.hysteresis, cfg.min_loss_scale)
        self.schedule = LRSchedule(cfg.sched_type, cfg.sched_params, cfg.lr)
        self.timers = PhaseTimers(cfg.wall_clock_breakdown, self.on_cuda)
        self.comm_stream = torch.cuda.Stream(device=self.device) if self.on_cuda else None
        self.comm_mode = comm
        self._loss_accum = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.last_metrics: Dict[str, float] = {}
        self.fault_hooks: Dict[str, Callable] = {}
        self.fused = None
        self._symm: Dict[str, Any] = {}
        # "auto" prefers the fused peer-memory kernels (the product) and falls back to NCCL (the baseline)
        # when peers cannot be mapped (no P2P/IPC) or the dtype is not bf16
        self._want_fused = (comm in ("fused", "auto") and self.on_cuda and self.world_size > 1
                            and self.cfg.reduce_scatter                  # reduce_scatter: false asks for all-reduce
                            and self.cfg.compute_dtype == torch.bfloat16
                            and self.cfg.offload_optimizer == "none"     # K3 updates device-resident state
                            and self.cfg.offload_param == "none"         # K2 / pull gather read device-resident shards
                            and os.environ.get("B200ZERO_COMM", comm) != "nccl")
        if comm in ("fused", "auto") and self.on_cuda and self.world_size > 1 and not self._want_fused:
            why = ("reduce_scatter is false" if not self.cfg.reduce_scatter else
                   "compute dtype is not bf16" if self.cfg.compute_dtype != torch.bfloat16 else
                   "optimizer state is offloaded" if self.cfg.offload_optimizer != "none" else
                   "parameters are offloaded" if self.cfg.offload_param != "none" else "B200ZERO_COMM=nccl")
            _log.warning("fused peer-memory collectives declined (%s): this job runs on NCCL collectives", why)
        if self._want_fused and comm == "auto":
            self._want_fused = self._probe_peer_memory()
            if self._want_fused:
                comm = self.comm_mode = "fused"
        self._in_backward = False
        self._reduce_now = True
        # Set by whoever learns that this rank is about to lose its machine (SIGTERM handler, spot probe). The bit rides
        # in the per-step stats all-reduce, so every rank sees the same answer at the same step boundary.
        self.preempt_requested = False
        hb_dir = os.environ.get("B200ZERO_HEARTBEAT_DIR")
        self._heartbeat_path = os.path.join(hb_dir, f"rank{self.rank}") if hb_dir else None
        self._prefetch_evt: Dict[int, Any] = {}

        # ---- parameters, units, layout ----
        self.source = module if isinstance(module, ParamSource) else None
        if self.source is not None:
            self.module = None
            self.blocks = []
            unit_names = list(self.source.unit_names)
            specs = list(self.source.specs())
            self.named_params = []
            self._param_of = {}
        else:
            module.to(self.device)
            unit_names, self.blocks, unit_of = discover_units(module)
            seen, plist = set(), []
            for name, p in module.named_parameters():
                if id(p) in seen:
                    continue
                seen.add(id(p))
                plist.append((name, p, unit_of.get(id(p), 0)))
            self.named_params = [(n, p) for n, p, _ in plist]
            specs = [(n, tuple(p.shape), u) for n, p, u in plist]
            self._param_of = {n: p for n, p, _ in plist}
            if getattr(module, "activation_checkpointing", None) is not None:
                module.activation_checkpointing = cfg.act_ckpt
            if broadcast_init and self.world_size > 1:
                for _, p in self.named_params:  # X7: identical start on every rank
                    dist.broadcast(p.data, src=0)
        self.layout: Layout = build_layout(specs, unit_names, self.world_size, partition=self.stage >= 1,
                                           persistence_threshold=cfg.persistence_threshold)
        self._build_buffers()
        if self.source is None:
            self._install_hooks()
        if self.comm_mode == "auto":
            self.comm_mode = "nccl" if self.on_cuda else "gloo"
        if self.comm_mode == "fused":
            if self._want_fused:
                from ..parallel.fused import FusedComm
                self.fused = FusedComm(self)
            else:
                self.comm_mode = "nccl" if self.on_cuda else "gloo"  # single rank / CPU: nothing to fuse
        from .config import audit
        self.config_notes = audit(self.cfg, [u.full_numel for u in self.layout.units], self.world_size,
                                  fused=self.fused is not None)
        if self.rank == 0:
            for note in self.config_notes:
                _log.warning("config: %s", note)

    # ------------------------------------------------------------------ buffers
    def _slice_rows(self, spec: ParamSpec, full: torch.Tensor) -> torch.Tensor:
        if not spec.partitioned:
            return full.reshape(-1)
        rows = spec.shape[0] // self.world_size
        return full[self.rank * rows:(self.rank + 1) * rows].reshape(-1)

    def _probe_peer_memory(self) -> bool:
        """Can every rank map every other rank's memory (CUDA IPC over NVLink)? Collective; all ranks
        return the same answer."""
        ok = 1
        try:
            from ..parallel.symm import SymmetricBuffer
            probe = SymmetricBuffer(4096, device=self.device)
            probe.close()
        except Exception:
            ok = 0
        t = torch.tensor([ok], device=self.device, dtype=torch.int32)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    def _zeros(self, numel: int, dtype: torch.dtype, name: str, symmetric: bool = False) -> torch.Tensor:
        """Device buffer; peer-mapped (CUDA IPC) when the fused backend will touch it from other ranks."""
        if symmetric and self._want_fused:
            from ..parallel.fused import FusedComm
            return FusedComm.alloc(self, numel, dtype, name)
        return torch.zeros(numel, dtype=dtype, device=self.device)

    def _build_buffers(self) -> None:
        L, cfg, dev, cd = self.layout, self.cfg, self.device, self.compute_dtype
        state_dev = torch.device("cpu") if cfg.offload_optimizer in ("cpu", "nvme") else dev
        pin = state_dev.type == "cpu" and self.on_cuda and cfg.pin_memory
        self._nvme = None
        if cfg.offload_optimizer == "nvme":
            # ZeRO-Infinity tier: state lives in files, host RAM only holds the staging slots (engine/nvme.py)
            from .nvme import NvmeStateStore, default_nvme_path
            self._nvme = NvmeStateStore(cfg.nvme_path or default_nvme_path(), self.rank, L.shard_numel,
                                        buffer_count=cfg.offload_buffer_count, pin=pin, grad_dtype=cfg.accum_dtype,
                                        chunk=int(os.environ.get("B200ZERO_NVME_CHUNK", 1 << 24)))
            self.master, self.exp_avg, self.exp_avg_sq = (self._nvme.view(n) for n in ("master", "exp_avg", "exp_avg_sq"))
            self.exp_avg.zero_()
            self.exp_avg_sq.zero_()
        else:
            mk = lambda: torch.zeros(L.shard_numel, dtype=torch.float32, device=state_dev, pin_memory=pin)  # noqa: E731
            self.master, self.exp_avg, self.exp_avg_sq = mk(), mk(), mk()
        self.state_device = state_dev
        self.shard_grads = torch.zeros(L.shard_numel, dtype=cfg.accum_dtype, device=dev)
        # offload_param (ref deepspeed_launcher.py:40,205-212; on in the reference's default config and its 70b preset):
        # the rank's 16-bit parameter shard lives in pinned host memory; a unit's rows are copied to a small device
        # staging buffer on the stream that gathers the unit (the prefetch stream when overlap_comm is on), so HBM
        # holds gathered blocks only. "nvme" is served from the same host tier.
        self.param_offload = cfg.offload_param in ("cpu", "nvme") and self.stage >= 3 and len(L.units) - 1 >= 2
        if cfg.offload_param in ("cpu", "nvme") and not self.param_offload:
            _log.warning("offload_param=%s ignored: it needs ZeRO-3 and a model with at least two blocks", cfg.offload_param)
        if cfg.offload_param == "nvme" and self.param_offload:
            _log.warning("offload_param=nvme: parameter shards are kept in pinned host memory (the NVMe tier holds "
                         "optimizer state only)")
        if self.param_offload:
            self.shard_params = torch.zeros(L.shard_numel, dtype=cd, device="cpu",
                                            pin_memory=self.on_cuda and cfg.pin_memory)
            self._pstage: Dict[Any, torch.Tensor] = {}
        else:
            self.shard_params = self._zeros(L.shard_numel, cd, "shard_params", symmetric=self.stage >= 3)
        # master <- initial fp32 values, own rows only
        for u in L.units:
            for spec in u.params:
                dst = self.master[spec.shard_offset:spec.shard_offset + spec.shard_numel]
                if self.source is not None:
                    rows = spec.shape[0] // self.world_size if spec.partitioned else spec.shape[0]
                    r0 = self.rank * rows if spec.partitioned else 0
                    dst.copy_(self.source.init_rows(spec.name, spec.shape, r0, r0 + rows).reshape(-1))
                else:
                    dst.copy_(self._slice_rows(spec, self._param_of[spec.name].data.detach().float()))
        self._refresh_shard_params()
        # parameter / gradient buffers
        nblk = len(L.units) - 1
        self._n_pslots = 2
        # One rank: nothing to gather or reduce — parameters and gradients *are* the shard buffers
        # (per unit, the shard layout of replicated tensors equals the unit's full layout).
        self._single = self.world_size == 1 and L.part_numel == 0 and not self.param_offload
        alias_grads = self._single and cfg.accum_dtype == cd
        if self._single:
            self._unit_pbuf = [self.shard_params[u.repl_shard_start:u.repl_shard_end] for u in L.units]
            self._resident = [True] * len(L.units)
            self._slot_owner = [None] * self._n_pslots
        elif self.stage >= 3 and nblk >= 2:
            blk_numel = max(u.full_numel for u in L.units[1:])
            # stage3_max_live_parameters (ref deepspeed_launcher.py:188) bounds how many gathered blocks stay alive:
            # at least the two the prefetch pipeline needs, at most all of them. Blocks still alive when backward
            # turns around (stage3_max_reuse_distance, :189: reuse is that close) are not gathered a second time.
            self._n_pslots = max(2, min(nblk, cfg.max_live_parameters // max(blk_numel, 1)))
            if cfg.max_reuse_distance < blk_numel:
                self._n_pslots = 2               # nothing is worth keeping beyond the pipeline's own two buffers
            self._n_pslots = int(os.environ.get("B200ZERO_PARAM_SLOTS", self._n_pslots))
            # symmetric (with the fused backend): the optimizer kernel publishes the root unit's new rows into it
            self._pbuf_root = self._zeros(L.units[0].full_numel, cd, "pbuf_root", symmetric=True)
            self._pbuf_blk = [torch.zeros(blk_numel, dtype=cd, device=dev) for _ in range(self._n_pslots)]
            self._unit_pbuf = [self._pbuf_root] + [self._pbuf_blk[self._pslot(i)] for i in range(1, len(L.units))]
            self._resident = [False] * len(L.units)      # which unit's weights currently sit in its buffer
            self._slot_owner: List[Optional[int]] = [None] * self._n_pslots
        else:
            total = sum(u.full_numel for u in L.units)
            self._pbuf_all = self._zeros(total, cd, "pbuf_all", symmetric=True)
            off, self._unit_pbuf = 0, []
            for u in L.units:
                self._unit_pbuf.append(self._pbuf_all[off:off + u.full_numel])
                off += u.full_numel
            self._resident = [True] * len(L.units)
            self._slot_owner = [None] * self._n_pslots
        self.pooled_params = self.stage >= 3 and nblk >= 2 and not self._single
        self.pooled_grads = self.stage >= 2 and nblk >= 2 and not alias_grads
        self._grads_alias = alias_grads
        if alias_grads:
            self._unit_gbuf = [self.shard_grads[u.repl_shard_start:u.repl_shard_end] for u in L.units]
        elif self.pooled_grads:
            blk_numel = max(u.full_numel for u in L.units[1:])
            self._gbuf_root = self._zeros(L.units[0].full_numel, cd, "gbuf_root", symmetric=True)
            self._gbuf_blk = [self._zeros(blk_numel, cd, f"gbuf_blk{i}", symmetric=True) for i in range(2)]
            self._unit_gbuf = [self._gbuf_root] + [self._gbuf_blk[(i - 1) % 2] for i in range(1, len(L.units))]
        else:
            total = sum(u.full_numel for u in L.units)
            self._gbuf_all = self._zeros(total, cd, "gbuf_all", symmetric=True)
            off, self._unit_gbuf = 0, []
            for u in L.units:
                self._unit_gbuf.append(self._gbuf_all[off:off + u.full_numel])
                off += u.full_numel
        self._gslot_free_evt: List[Optional[Any]] = [None, None]  # grad pool slot reusable after this event
        # bind param.data / param.grad views (once); a ParamSource engine reads views on demand
        if self.source is None:
            for u in L.units:
                pb, gb = self._unit_pbuf[u.index], self._unit_gbuf[u.index]
                for spec in u.params:
                    p = self._param_of[spec.name]
                    view = pb[spec.full_offset:spec.full_offset + spec.numel].view(spec.shape)
                    if self._resident[u.index]:
                        view.copy_(p.data.to(cd))
                    p.data = view
                    p.grad = gb[spec.full_offset:spec.full_offset + spec.numel].view(spec.shape)
        elif not self.pooled_params:
            for ui in range(len(L.units)):
                self._gather_unit(ui)
        # small scalars used by step()
        self._stats = torch.zeros(4, dtype=torch.float32, device=dev)      # [sumsq_part, nonfinite, loss, sumsq_repl]
        self._flag = torch.zeros(1, dtype=torch.int32, device=dev)
        self._coef = torch.ones(1, dtype=torch.float32, device=dev)
        self._norm = torch.zeros(1, dtype=torch.float32, device=dev)
        self._sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        if self.on_cuda:
            self._stats_host = torch.zeros(5, dtype=torch.float32).pin_memory()
        if self.stage >= 3 and self.pooled_params:
            self._gather_unit(0)

    # ------------------------------------------------------------------ hooks
    def _install_hooks(self) -> None:
        L = self.layout
        self._pending = [0] * len(L.units)
        self._need = [sum(1 for s in u.params if self._param_of[s.name].requires_grad) for u in L.units]
        self._reduced = [False] * len(L.units)
        for u in L.units:
            for spec in u.params:
                p = self._param_of[spec.name]
                if p.requires_grad:
                    p.register_post_accumulate_grad_hook(self._make_grad_hook(u.index))
        for bi, blk in enumerate(self.blocks):
            ui = bi + 1
            blk.register_forward_pre_hook(self._make_pre_fwd(ui))
            blk.register_full_backward_pre_hook(self._make_pre_bwd(ui))

    def _make_grad_hook(self, ui: int):
        def hook(_p):
            self._pending[ui] -= 1
            if self._pending[ui] == 0 and self._reduce_now:
                self._reduce_unit(ui)
        return hook

    def _make_pre_fwd(self, ui: int):
        def hook(_m, _inp):
            if self.pooled_params:
                self._ensure_resident(ui)
                nxt = ui + 1 if not self._in_backward else ui - 1
                if 1 <= nxt < len(self.layout.units):
                    self._prefetch(nxt)
        return hook

    def _make_pre_bwd(self, ui: int):
        def hook(_m, _gout):
            if self.pooled_params:
                self._ensure_resident(ui)
                if ui - 1 >= 1:
                    self._prefetch(ui - 1)
            if self.pooled_grads:
                evt = self._gslot_free_evt[(ui - 1) % 2]
                if evt is not None and self.on_cuda:
                    torch.cuda.current_stream().wait_event(evt)
        return hook

    # ------------------------------------------------------------------ stage-3 parameter residency
    def _pslot(self, ui: int) -> int:
        """Which of the gathered-block buffers unit ``ui`` (>= 1) lives in."""
        return (ui - 1) % self._n_pslots

    def _gather_unit(self, ui: int, stream=None) -> None:
        """All-gather unit ``ui``'s rows into its buffer (X2). Replicated tensors are local copies."""
        if self._single:
            return                      # parameters are views of the shard: always current
        u = self.layout.units[ui]
        pb = self._unit_pbuf[ui]
        ctx = torch.cuda.stream(stream) if (stream is not None and self.on_cuda) else nullcontext()
        with ctx:
            if self.fused is not None and self.fused.gather_unit(ui):
                pass
            else:
                shard, base_p, base_r = self.shard_params, 0, 0
                if self.param_offload:
                    shard, base_p, base_r = self._stage_unit_shard(ui)
                for spec in u.params:
                    dst = pb[spec.full_offset:spec.full_offset + spec.numel]
                    o = spec.shard_offset - (base_p if spec.partitioned else base_r)
                    src = shard[o:o + spec.shard_numel]
                    if spec.partitioned:
                        _all_gather(dst, src)
                    else:
                        dst.copy_(src)
        self._resident[ui] = True
        if ui >= 1 and self.pooled_params:
            slot = self._pslot(ui)
            prev = self._slot_owner[slot]
            if prev is not None and prev != ui:
                self._resident[prev] = False
            self._slot_owner[slot] = ui

    def _stage_unit_shard(self, ui: int):
        """Parameter offload: bring unit ``ui``'s rows of the host-resident shard onto the device (two contiguous
        pinned copies on the current stream: the partitioned range and the replicated range). Returns the staging
        tensor and the shard offsets its two halves start from."""
        u = self.layout.units[ui]
        n_part, n_repl = u.part_shard_end - u.part_shard_start, u.repl_shard_end - u.repl_shard_start
        key = "root" if ui == 0 else self._pslot(ui)
        stg = self._pstage.get(key)
        if stg is None or stg.numel() < n_part + n_repl:
            stg = torch.empty(n_part + n_repl, dtype=self.compute_dtype, device=self.device)
            self._pstage[key] = stg
        if n_part:
            stg[:n_part].copy_(self.shard_params[u.part_shard_start:u.part_shard_end], non_blocking=True)
        if n_repl:
            stg[n_part:n_part + n_repl].copy_(self.shard_params[u.repl_shard_start:u.repl_shard_end], non_blocking=True)
        self.h2d_param_bytes = getattr(self, "h2d_param_bytes", 0) + (n_part + n_repl) * stg.element_size()
        # a replicated tensor's offset inside the staging buffer = its shard offset - repl_start + n_part
        return stg, u.part_shard_start, u.repl_shard_start - n_part

    def _refresh_shard_params(self, lo: int = 0, hi: Optional[int] = None) -> None:
        """16-bit shard <- fp32 master for elements [lo, hi) (wherever either of them lives), in bounded chunks."""
        hi = self.layout.shard_numel if hi is None else hi
        step = 1 << 26
        for a in range(lo, hi, step):
            b = min(a + step, hi)
            src = self.master[a:b]
            if self.shard_params.is_cuda or not src.is_cuda:
                self.shard_params[a:b].copy_(src.to(self.shard_params.device).to(self.compute_dtype))
            else:
                self.shard_params[a:b].copy_(src.to(self.compute_dtype))       # cast on the device, 2 bytes/elem over PCIe

    def _ensure_resident(self, ui: int) -> None:
        if self._resident[ui] and (ui == 0 or self._slot_owner[self._pslot(ui)] == ui):
            evt = self._prefetch_evt.pop(ui, None)
            if evt is not None:
                torch.cuda.current_stream().wait_event(evt)
            return
        self._gather_unit(ui)

    def _prefetch(self, ui: int) -> None:
        """Issue the gather of unit ``ui`` on the comm stream (``stage3_prefetch_bucket_size`` > 0)."""
        if self._resident[ui] and self._slot_owner[self._pslot(ui)] == ui:
            return
        if not self.on_cuda or not self.cfg.overlap_comm or self.cfg.prefetch_bucket_size <= 0:
            return  # gathered synchronously on demand
        # the slot's previous owner (ui±2) finished executing before the block that triggered us started
        self.comm_stream.wait_stream(torch.cuda.current_stream())
        self._gather_unit(ui, stream=self.comm_stream)
        evt = torch.cuda.Event()
        evt.record(self.comm_stream)
        self._prefetch_evt[ui] = evt

    # ------------------------------------------------------------------ gradient reduction
    def _reduce_unit(self, ui: int) -> None:
        """Reduce-scatter (X1) / all-reduce (X4) unit ``ui``'s gradients into the rank's shard."""
        if self._reduced[ui]:
            return
        self._reduced[ui] = True
        hook = self.fault_hooks.get("pre_reduce")
        if hook is not None:
            hook(self, ui)
        if self._grads_alias:
            return                      # single rank: the unit's gradients already live in the shard
        u = self.layout.units[ui]
        gb = self._unit_gbuf[ui]
        use_side = self.on_cuda and self.cfg.overlap_comm
        if use_side:
            self.comm_stream.wait_stream(torch.cuda.current_stream())
        ctx = torch.cuda.stream(self.comm_stream) if use_side else nullcontext()
        scale = 1.0 / self.world_size  # mean over data-parallel ranks
        with ctx, self.timers.phase("reduce"):
            if self.fused is not None and self.fused.reduce_unit(ui, scale):
                pass
            else:
                rdt = self.cfg.reduce_dtype
                # prescale_gradients / gradient_predivide_factor (ref deepspeed_launcher.py:141-142 emits the keys):
                # divide before the low-precision sum so it cannot overflow, finish the mean afterwards
                pre = float(self.cfg.predivide_factor) if self.cfg.prescale_gradients else 1.0
                if self.cfg.prescale_gradients and pre == 1.0:
                    pre = float(self.world_size)
                post = scale * pre
                for spec in u.params:
                    g = gb[spec.full_offset:spec.full_offset + spec.numel]
                    dst = self.shard_grads[spec.shard_offset:spec.shard_offset + spec.shard_numel]
                    if self.world_size == 1:
                        dst.add_(g.to(dst.dtype))
                        continue
                    gr = g if g.dtype == rdt else g.to(rdt)
                    if pre != 1.0:
                        gr = gr / pre
                    if spec.partitioned and self.cfg.reduce_scatter:
                        dst.add_(_reduce_scatter(gr, self.rank, self.world_size).to(dst.dtype), alpha=post)
                    elif spec.partitioned:            # reduce_scatter: false -> all-reduce, then keep my rows
                        full = gr.clone() if gr.data_ptr() == g.data_ptr() else gr
                        dist.all_reduce(full)
                        n = spec.shard_numel
                        dst.add_(full[self.rank * n:(self.rank + 1) * n].to(dst.dtype), alpha=post)
                    else:
                        if gr.data_ptr() == g.data_ptr():
                            gr = gr.clone()
                        dist.all_reduce(gr)
                        dst.add_(gr.to(dst.dtype), alpha=post)
            if getattr(self, "_zero_after_reduce", True):
                gb.zero_()         # autograd accumulates in place; the native executor overwrites instead
            if use_side and self.pooled_grads and ui >= 1:
                evt = torch.cuda.Event()
                evt.record(self.comm_stream)
                self._gslot_free_evt[(ui - 1) % 2] = evt

    # ------------------------------------------------------------------ public protocol
    def __call__(self, *args, **kwargs):
        return self.forward(*args, **kwargs)

    def train(self, mode: bool = True):
        self.module.train(mode)
        return self

    def eval(self):
        return self.train(False)

    def is_gradient_accumulation_boundary(self) -> bool:
        return (self.micro_steps + 1) % self.cfg.grad_accum == 0

    def forward(self, *args, **kwargs):
        self._in_backward = False
        if self.pooled_params:
            self._ensure_resident(0)
        with self.timers.phase("forward"):
            return self.module(*args, **kwargs)

    @torch.no_grad()
    def eval_loss(self, *args, **kwargs) -> torch.Tensor:
        """Validation loss of one batch: the module's forward under ``no_grad`` (no gradient hooks fire)."""
        was_training = self.module.training
        self.module.eval()
        try:
            return self.forward(*args, **kwargs)
        finally:
            self.module.train(was_training)

    def backward(self, loss: torch.Tensor) -> None:
        cfg = self.cfg
        self._in_backward = True
        boundary = self.is_gradient_accumulation_boundary()
        self._reduce_now = self.stage >= 2 or boundary
        self._pending = list(self._need)
        self._reduced = [False] * len(self.layout.units)
        if self.fused is not None:
            self.fused.begin_micro_step()
        hook = self.fault_hooks.get("pre_backward")
        if hook is not None:
            loss = hook(self, loss)
        self._loss_accum += loss.detach().float() / cfg.grad_accum
        with self.timers.phase("backward"):
            (loss * (self.scaler.scale / cfg.grad_accum)).backward()
        if self._reduce_now:
            for ui in range(len(self.layout.units) - 1, -1, -1):  # units whose hooks never all fired
                if not self._reduced[ui]:
                    self._reduce_unit(ui)
        if self.on_cuda and cfg.overlap_comm:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        hook = self.fault_hooks.get("post_backward")
        if hook is not None:
            hook(self)
        self._in_backward = False

    def step(self) -> Optional[Dict[str, float]]:
        """Call after every ``backward``; does the optimizer step at accumulation boundaries.
        Returns the step's metrics dict at a boundary, else None."""
        self.micro_steps += 1
        self.global_samples += self.cfg.micro_batch * self.world_size
        if self.micro_steps % self.cfg.grad_accum != 0:
            return None
        return self._optimizer_step()

    # ------------------------------------------------------------------ optimizer step
    def _optimizer_step(self) -> Dict[str, float]:
        cfg, L = self.cfg, self.layout
        with self.timers.phase("step"):
            st = self._stats
            st.zero_()
            if self.fused is not None:       # Σg² and the flag came out of the reduce-scatter kernels
                st[0] = self.fused.sumsq_part[0]
                st[3] = self.fused.sumsq_repl[0]
            else:
                if L.part_numel:
                    O.sumsq(self.shard_grads[:L.part_numel], st[0:1], nonfinite_flag=self._flag)
                if L.shard_numel > L.part_numel:
                    O.sumsq(self.shard_grads[L.part_numel:], st[3:4], nonfinite_flag=self._flag)
            st[1] = self._flag.float()[0]
            st[2] = self._loss_accum[0] / self.world_size
            want_stop = 1.0 if self.preempt_requested else 0.0
            if self.world_size > 1:
                repl = st[3].clone()
                st[3] = want_stop             # fourth slot: "some rank was told to stop" (summed, so > 0 means yes)
                if self.fused is None or not self.fused.allreduce_stats(st):
                    dist.all_reduce(st)       # Σ‖g‖² (X6), overflow flag (X5), mean loss (X8) in one message
                stop_votes = st[3].clone()
                st[3] = repl
            else:
                stop_votes = torch.full((), want_stop, dtype=torch.float32, device=self.device)
            self._sumsq[0] = st[0] + st[3]
            self._flag[0] = (st[1] > 0).int()
            inv_scale = 1.0 / self.scaler.scale
            O.clip_coef(self._sumsq, cfg.grad_clip * self.scaler.scale if cfg.grad_clip > 0 else 0.0, inv_scale,
                        self._coef, self._norm, self._flag)
            lr = self.schedule.lr_at(self.global_steps)
            step_no = self.global_steps - self.skipped_steps + 1
            if self.fused is not None:
                self.fused.adam_and_publish(lr, step_no)
            elif self._nvme is not None:
                from .nvme import nvme_adam_step
                nvme_adam_step(self, lr, step_no)
            elif self.state_device.type == "cpu" and self.on_cuda:
                self._offloaded_adam(lr, step_no)
            else:
                O.adamw_step(self.master, self.exp_avg, self.exp_avg_sq, self.shard_grads, lr=lr, beta1=cfg.betas[0],
                             beta2=cfg.betas[1], eps=cfg.eps, weight_decay=cfg.weight_decay, step=step_no,
                             grad_scale_dev=self._coef, skip_flag=self._flag,
                             param_out=self.shard_params if self.compute_dtype == torch.bfloat16 and self.on_cuda
                             and not self.param_offload else None)
                if not (self.compute_dtype == torch.bfloat16 and self.on_cuda) or self.param_offload:
                    self._refresh_shard_params()
            # the fused Adam kernel zeroes the shard as it reads it; a single-rank executor whose producers overwrite
            # every gradient in place (gradients alias the shard) needs no zeroing at all
            overwritten = self._grads_alias and not getattr(self, "_zero_after_reduce", True)
            if self.fused is None and not overwritten:
                self.shard_grads.zero_()
            # host read: the one sync of the step (feeds the loss monitor)
            if self.on_cuda:
                self._stats_host.copy_(torch.stack([self._sumsq[0], self._flag[0].float(), st[2], self._norm[0] * inv_scale,
                                                    stop_votes]), non_blocking=True)
                torch.cuda.current_stream().synchronize()
                vals = self._stats_host.tolist()
            else:
                vals = [float(self._sumsq[0]), float(self._flag[0]), float(st[2]), float(self._norm[0]) * inv_scale,
                        float(stop_votes)]
            self._flag.zero_()                 # clean for the next step's reduce / norm kernels
            overflow = vals[1] > 0
            self.scaler.update(overflow)
            if overflow:
                self.skipped_steps += 1
            # updated rows -> everyone (X3); stage 3 re-gathers lazily instead
            with self.timers.phase("allgather"):
                if self.stage >= 3 and self.pooled_params:
                    self._resident = [False] * len(L.units)
                    self._slot_owner = [None] * self._n_pslots
                    if self.fused is not None and self.fused.can_push_root():
                        self._resident[0] = True     # the optimizer kernel pushed the root unit's rows everywhere
                    else:
                        self._gather_unit(0)
                elif self.fused is not None:
                    pass                        # rows were pushed into every rank's buffers by the Adam kernel
                else:
                    for ui in range(len(L.units)):
                        self._gather_unit(ui)
            self._loss_accum.zero_()
        self.global_steps += 1
        self.schedule.last_step = self.global_steps
        if self._heartbeat_path:                 # the launcher's stall detection watches this file (engine/launch.py)
            try:
                with open(self._heartbeat_path, "w") as f:
                    f.write(str(self.global_steps))
            except OSError:
                pass
        m = {"step": self.global_steps, "loss": vals[2], "grad_norm": vals[3], "grad_norm_sq": vals[3] ** 2,
             "nonfinite": overflow, "lr": lr, "loss_scale": self.scaler.scale, "skipped": overflow,
             "preempt": vals[4] > 0}
        if cfg.wall_clock_breakdown:
            m["timers_ms"] = self.timers.collect()
        self.last_metrics = m
        return m

    def _offloaded_adam(self, lr: float, step_no: int) -> None:
        from .offload import host_adam_step
        host_adam_step(self, lr, step_no)

    # ------------------------------------------------------------------ views for module-less engines
    def param_view(self, name: str) -> torch.Tensor:
        spec = self.layout.by_name[name]
        return self._unit_pbuf[spec.unit][spec.full_offset:spec.full_offset + spec.numel].view(spec.shape)

    def grad_view(self, name: str) -> torch.Tensor:
        spec = self.layout.by_name[name]
        return self._unit_gbuf[spec.unit][spec.full_offset:spec.full_offset + spec.numel].view(spec.shape)

    # ------------------------------------------------------------------ introspection / state
    def get_lr(self) -> List[float]:
        return [self.schedule.lr_at(self.global_steps)]

    def memory_report(self) -> Dict[str, float]:
        gib = float(1 << 30)
        L = self.layout
        pbufs = {b.data_ptr(): b for b in self._unit_pbuf}
        gbufs = {b.data_ptr(): b for b in self._unit_gbuf}
        return {"params_total": L.total_params, "shard_numel": L.shard_numel,
                "optimizer_state_gib": 3 * L.shard_numel * 4 / gib,
                "grad_shard_gib": L.shard_numel * self.shard_grads.element_size() / gib,
                "param_buffers_gib": sum(b.numel() * b.element_size() for b in pbufs.values()) / gib,
                "grad_buffers_gib": sum(b.numel() * b.element_size() for b in gbufs.values()) / gib}

    def full_param(self, name: str) -> torch.Tensor:
        """Gathered fp32 master copy of one parameter (collective when partitioned)."""
        spec = self.layout.by_name[name]
        local = self.master[spec.shard_offset:spec.shard_offset + spec.shard_numel].to(self.device)
        if not spec.partitioned:
            return local.view(spec.shape).clone()
        out = torch.empty(spec.numel, dtype=torch.float32, device=self.device)
        _all_gather(out, local.contiguous())
        return out.view(spec.shape)

    def full_state_dict(self) -> Dict[str, torch.Tensor]:
        """Every parameter as a full fp32 tensor from the master copy, on the CPU (collective: all ranks call it;
        all ranks receive it). Feed it to ``models.hf.to_hf_state_dict`` to leave for the transformers layout."""
        return {spec.name: self.full_param(spec.name).cpu() for u in self.layout.units for spec in u.params}

    def module_state_dict(self) -> Dict[str, torch.Tensor]:
        """16-bit (compute dtype) weights by name, gathered (``stage3_gather_16bit_weights_on_model_save``)."""
        sd = {}
        for u in self.layout.units:
            if not self._resident[u.index]:
                self._gather_unit(u.index)
            for spec in u.params:
                sd[spec.name] = self.param_view(spec.name).detach().clone().cpu()
        return sd

    def load_full_params(self, named: Dict[str, torch.Tensor], strict: bool = True) -> List[str]:
        """Initialise the model from full (unsharded) tensors by parameter name — e.g. a converted Hugging Face
        checkpoint (``models/hf.py``) or the output of ``zero_to_fp32.py``. Every rank passes the same dict and keeps
        its own rows; Adam moments and step counters are left untouched. Returns the names that were not supplied."""
        missing = []
        for u in self.layout.units:
            for spec in u.params:
                t = named.get(spec.name)
                if t is None:
                    missing.append(spec.name)
                    continue
                if tuple(t.shape) != tuple(spec.shape):
                    raise ValueError(f"{spec.name}: expected shape {tuple(spec.shape)}, got {tuple(t.shape)}")
                dst = self.master[spec.shard_offset:spec.shard_offset + spec.shard_numel]
                dst.copy_(self._slice_rows(spec, t.detach().to(torch.float32)))
        if strict and missing:
            raise KeyError(f"parameters not supplied: {missing[:8]}{' ...' if len(missing) > 8 else ''}")
        self._refresh_shard_params()
        if _dist_ready() and self.world_size > 1:
            if self.on_cuda:
                torch.cuda.synchronize(self.device)      # peers read this shard directly in the fused gather
            dist.barrier()
        if self.stage >= 3 and self.pooled_params:
            self._resident = [False] * len(self.layout.units)
            self._slot_owner = [None] * self._n_pslots
            self._gather_unit(0)
        else:
            for ui in range(len(self.layout.units)):
                self._gather_unit(ui)
        if _dist_ready() and self.world_size > 1:
            dist.barrier()
        return missing

    def save_checkpoint(self, save_dir: str, tag: Optional[str] = None, client_state: Optional[Dict] = None,
                        async_write: bool = False):
        from .checkpoint import save_checkpoint
        return save_checkpoint(self, save_dir, tag, client_state or {}, async_write=async_write)

    def load_checkpoint(self, load_dir: str, tag: Optional[str] = None):
        from .checkpoint import load_checkpoint
        self._prefetch_evt.clear()
        return load_checkpoint(self, load_dir, tag)


# ---------------------------------------------------------------------- collectives with gloo fallbacks
def _all_gather(dst_flat: torch.Tensor, src_flat: torch.Tensor) -> None:
    if not _dist_ready() or dist.get_world_size() == 1:
        dst_flat.copy_(src_flat)
        return
    try:
        dist.all_gather_into_tensor(dst_flat, src_flat.contiguous())
    except (RuntimeError, NotImplementedError):
        ws = dist.get_world_size()
        parts = list(dst_flat.view(ws, -1).unbind(0))
        dist.all_gather(parts, src_flat.contiguous())


def _reduce_scatter(full_flat: torch.Tensor, rank: int, ws: int) -> torch.Tensor:
    n = full_flat.numel() // ws
    if full_flat.is_cuda:
        out = torch.empty(n, dtype=full_flat.dtype, device=full_flat.device)
        dist.reduce_scatter_tensor(out, full_flat.contiguous())
        return out
    tmp = full_flat.clone()  # gloo has no reduce-scatter: all-reduce then slice
    dist.all_reduce(tmp)
    return tmp[rank * n:(rank + 1) * n]


def initialize(model: nn.Module, config: Union[str, Dict[str, Any], EngineConfig], **kw) -> ZeroEngine:
    """DeepSpeed-style constructor: ``engine = initialize(model=..., config=...)``."""
    return ZeroEngine(model, config, **kw)
