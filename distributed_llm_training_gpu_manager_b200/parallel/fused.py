"""Fused-collective backend for :class:`ZeroEngine` (``comm="fused"``): the per-unit gradient
reduce-scatter (K1) and the optimizer step with push all-gather (K3) run as our own kernels over
peer-mapped memory instead of NCCL calls; ZeRO-3 weight gathers use the pull all-gather kernel
(and the all-gather→GEMM kernel K2 for the block's first GEMM, see ``ops/ag_gemm.py``).

Division of labour per optimizer step (cf. ``engine/zero.py`` which does the same with NCCL):

    backward, unit u done ──► bz_fused_reduce_scatter(u): pull my rows from all peers, fp32 sum in
                              rank order, ×1/ws, += shard, Σg² partials, non-finite flag
    after the last unit   ──► one 3-float NCCL all-reduce (Σ‖g‖², flag, loss) — 12 bytes, control only
    then                  ──► bz_fused_adam_push: AdamW on the shard (skip if flag; ×clip coef read
                              on device) and, for stage ≤ 2, store the new bf16 rows into every
                              peer's parameter buffer; epoch barrier so nobody reads early
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from ..engine.partition import _up
from ..ops import lib as _lib
from .symm import SymmetricBuffer, nvls_buffer_or_none

RS_CHANNEL, ADAM_CHANNEL, STATS_CHANNEL = 0, 1, 2


def seg_len(spec, world_size: int) -> int:
    """Element count a fused kernel walks for one tensor. The kernels move 8-element (16-byte) vectors, so a
    replicated tensor is handed over with its zero-filled padding (the layout pads every tensor to ``ALIGN``
    elements in the unit buffers and in the shard alike); a partitioned one must split into whole vectors."""
    if spec.partitioned:
        assert (spec.numel // world_size) % 8 == 0, f"{spec.name}: per-rank slice is not a whole number of vectors"
        return spec.numel
    return _up(spec.numel)


def segment_table(records: List[tuple], device) -> torch.Tensor:
    """records: (full_off, numel, shard_off, partitioned[, push]) -> int64 [n, 4] device tensor whose rows
    are the 32-byte Segment structs of csrc/fused_comm.cu."""
    # int32 partitioned + int32 push = one little-endian int64
    rows = [[r[0], r[1], r[2], int(bool(r[3])) | ((int(bool(r[4])) << 32) if len(r) > 4 else 0)] for r in records]
    assert all(r[1] % 8 == 0 for r in rows), "fused kernels walk whole 8-element vectors"
    return torch.tensor(rows, dtype=torch.int64, device=device).contiguous()


class FusedComm:
    def __init__(self, engine, rs_ctas: Optional[int] = None, adam_ctas: Optional[int] = None):
        import os
        self.push_root = False
        rs_ctas = rs_ctas or int(os.environ.get("B200ZERO_RS_CTAS", "128"))
        adam_ctas = adam_ctas or int(os.environ.get("B200ZERO_ADAM_CTAS", "148"))
        self.nvls_ctas = int(os.environ.get("B200ZERO_NVLS_CTAS", "32"))
        self.pull_version = int(os.environ.get("B200ZERO_PULL_AG", "1"))   # 2: division-free kernel (fused_comm_v2.cu)
        self.use_fused_gather = os.environ.get("B200ZERO_FUSED_GATHER", "1") == "1"
        self.use_fused_rs = os.environ.get("B200ZERO_FUSED_RS", "1") == "1"
        # 3: the "all ranks ready" wait of K1 is a one-warp launch in front of the reduce kernel (default);
        # 1: every CTA of the reduce kernel waits itself (B200ZERO_RS_SPLIT_WAIT=0)
        self.rs_barrier = 3 if os.environ.get("B200ZERO_RS_SPLIT_WAIT", "1") == "1" else 1
        assert engine.on_cuda, "fused communication needs GPUs"
        self.e = engine
        self.L = _lib.load(required=True)
        assert engine.world_size <= self.L.bz_comm_max_world()
        assert engine.compute_dtype == torch.bfloat16, "fused kernels reduce bf16 gradients"
        self.rs_ctas, self.adam_ctas = rs_ctas, adam_ctas
        dev = engine.device
        self.signal = SymmetricBuffer(self.L.bz_comm_signal_pad_bytes(), device=dev)
        self.workspace = torch.zeros(self.L.bz_fused_comm_workspace_bytes() // 4, dtype=torch.float32, device=dev)
        self.sumsq_part = torch.zeros(1, dtype=torch.float32, device=dev)
        self.sumsq_repl = torch.zeros(1, dtype=torch.float32, device=dev)
        self.epoch = {RS_CHANNEL: 0, ADAM_CHANNEL: 0, STATS_CHANNEL: 0}
        # X5/X6/X8 without NCCL (opt-in): the step's 4-float control vector is exchanged through the signal pads
        self.peer_stats = os.environ.get("B200ZERO_STATS_PEER", "0") == "1"
        self._sig_ptrs = self.signal.ptr_array()
        L = engine.layout
        # per-unit segment tables (offsets relative to the unit's buffer)
        self._unit_segs: Dict[int, torch.Tensor] = {}
        for u in L.units:
            self._unit_segs[u.index] = segment_table(
                [(s.full_offset, seg_len(s, engine.world_size), s.shard_offset, s.partitioned) for s in u.params], dev)
        self.launches = 0
        self.nvls_launches = 0

    # -- symmetric buffers are created by the engine through this hook --------------------------------
    @staticmethod
    def alloc(engine, numel: int, dtype: torch.dtype, name: str) -> torch.Tensor:
        import os
        es = torch.empty((), dtype=dtype).element_size()
        buf = None
        # NVLS (opt-in, experimental): gradient buffers additionally bound to an NVSwitch multicast object so
        # K1 reduces inside the switch. Everything else stays on the plain IPC mapping.
        if os.environ.get("B200ZERO_NVLS", "0") == "1" and name.startswith("gbuf"):
            buf = nvls_buffer_or_none(numel * es, engine.device)
        if buf is None:
            buf = SymmetricBuffer(numel * es, device=engine.device)
        engine._symm[name] = buf
        return buf.tensor(dtype, numel)

    def _ptrs_for(self, tensor: torch.Tensor):
        """Peer pointer array for a tensor living inside one of the engine's symmetric buffers."""
        p = tensor.data_ptr()
        for buf in self.e._symm.values():
            if buf.local_ptr <= p < buf.local_ptr + buf.nbytes:
                return buf.ptr_array(p - buf.local_ptr)
        raise RuntimeError("tensor is not in symmetric memory")

    def _mc_for(self, tensor: torch.Tensor) -> int:
        """Multicast address of the same bytes (0 when the buffer has no NVLS mapping)."""
        p = tensor.data_ptr()
        for buf in self.e._symm.values():
            if buf.local_ptr <= p < buf.local_ptr + buf.nbytes:
                mc = getattr(buf, "multicast_ptr", 0)
                return mc + (p - buf.local_ptr) if mc else 0
        return 0

    # -- K1 ------------------------------------------------------------------------------------
    def begin_micro_step(self) -> None:
        self.sumsq_part.zero_()
        self.sumsq_repl.zero_()

    def reduce_unit(self, ui: int, scale: float) -> bool:
        e = self.e
        if not self.use_fused_rs:
            return False
        gb = e._unit_gbuf[ui]
        self.epoch[RS_CHANNEL] += 1
        tail = (self._unit_segs[ui].data_ptr(), len(e.layout.units[ui].params),
                e.shard_grads.data_ptr(), int(e.shard_grads.dtype == torch.float32), scale, self.workspace.data_ptr(),
                self.sumsq_part.data_ptr(), self.sumsq_repl.data_ptr(), e._flag.data_ptr(), self.epoch[RS_CHANNEL],
                RS_CHANNEL, e.rank, e.world_size, self.rs_barrier, self.rs_ctas, _lib.stream())
        mc = self._mc_for(gb)
        if mc:
            tail = tail[:-2] + (self.nvls_ctas, tail[-1])      # the switch does the summing: fewer CTAs suffice
            rc = self.L.bz_fused_reduce_scatter_mc(self._ptrs_for(gb), self._sig_ptrs, mc, *tail)
            self.nvls_launches += 1
        else:
            rc = self.L.bz_fused_reduce_scatter(self._ptrs_for(gb), self._sig_ptrs, *tail)
        _lib.check(rc, "fused_reduce_scatter")
        _lib.add_launches(1)
        self.launches += 1
        return True

    # -- control exchange ------------------------------------------------------------------------
    def allreduce_stats(self, st: torch.Tensor) -> bool:
        """Sum the engine's 4-float stats vector over all ranks in place (rank-ordered, bit-identical everywhere).
        Returns False when the peer path is off (the caller then uses the NCCL all-reduce)."""
        if not self.peer_stats:
            return False
        assert st.numel() == 4 and st.dtype == torch.float32
        self.epoch[STATS_CHANNEL] += 1
        rc = self.L.bz_peer_allreduce4(self._sig_ptrs, st.data_ptr(), self.epoch[STATS_CHANNEL], STATS_CHANNEL,
                                       self.e.rank, self.e.world_size, _lib.stream())
        _lib.check(rc, "peer_allreduce4")
        return True

    # -- ZeRO-3 gather -------------------------------------------------------------------------
    def gather_unit(self, ui: int) -> bool:
        e = self.e
        if e.stage < 3 or "shard_params" not in e._symm or not self.use_fused_gather:
            return False
        fn = self.L.bz_pull_allgather_v2 if self.pull_version == 2 else self.L.bz_pull_allgather
        rc = fn(self._ptrs_for(e.shard_params), self._unit_segs[ui].data_ptr(),
                len(e.layout.units[ui].params), e._unit_pbuf[ui].data_ptr(), e.rank,
                e.world_size, self.rs_ctas, _lib.stream())
        _lib.check(rc, "pull_allgather")
        return True

    # -- K3 ------------------------------------------------------------------------------------
    def adam_and_publish(self, lr: float, step_no: int) -> None:
        e, cfg = self.e, self.e.cfg
        push = int(e.stage <= 2)
        self.epoch[ADAM_CHANNEL] += 1
        if push:
            if not hasattr(self, "_all_segs"):
                recs, off = [], 0
                for u in e.layout.units:
                    recs += [(off + s.full_offset, seg_len(s, e.world_size), s.shard_offset, s.partitioned)
                             for s in u.params]
                    off += u.full_numel
                self._all_segs = segment_table(recs, e.device)
                self._all_ptrs = self._ptrs_for(e._pbuf_all)
            segs, nseg, pptrs = self._all_segs, self._all_segs.shape[0], self._all_ptrs
        else:
            if not hasattr(self, "_all_segs"):
                # Stage 3: blocks are re-gathered when they run, but the root unit (embedding, final norm, LM head)
                # is needed by every step from its first kernel on. Its updated rows are published here, by the
                # optimizer kernel itself, into every rank's gathered root buffer (per-segment push flag) — there
                # is no separate gather of the root unit after the step.
                self.push_root = self.can_push_root()
                recs = [(s.full_offset if (self.push_root and u.index == 0) else 0, seg_len(s, e.world_size),
                         s.shard_offset, s.partitioned, self.push_root and u.index == 0)
                        for u in e.layout.units for s in u.params]
                self._all_segs = segment_table(recs, e.device)
                self._root_ptrs = self._ptrs_for(e._pbuf_root) if self.push_root else None
            segs, nseg, pptrs = self._all_segs, self._all_segs.shape[0], self._root_ptrs
        tail = (segs.data_ptr(), nseg, e.master.data_ptr(), e.exp_avg.data_ptr(),
                e.exp_avg_sq.data_ptr(), e.shard_grads.data_ptr(), int(e.shard_grads.dtype == torch.float32),
                e.shard_params.data_ptr(), lr, cfg.betas[0], cfg.betas[1], cfg.eps, cfg.weight_decay, step_no,
                e._coef.data_ptr(), e._flag.data_ptr(), self.workspace.data_ptr(), self.epoch[ADAM_CHANNEL],
                ADAM_CHANNEL, e.rank, e.world_size, push, 1, 1, self.adam_ctas, _lib.stream())
        rc = self.L.bz_fused_adam_push(pptrs, self._sig_ptrs, *tail)
        _lib.check(rc, "fused_adam_push")

    def debug_snapshot(self) -> dict:
        """What this rank's signal pad and host-side epoch counters look like right now, read on a private stream
        (works while the main / communication streams are stuck behind a kernel that waits for a peer). For every
        channel and phase: the epoch each rank has announced to us. Used by bench.py's stall watchdog."""
        self.prepare_debug_snapshot()
        side, host = self._dbg_stream, self._dbg_host
        words = host.numel()
        with torch.cuda.stream(side):
            host.copy_(self.signal.tensor(torch.int32, words), non_blocking=True)
        side.synchronize()
        ws, tab = self.e.world_size, host.tolist()
        names = {RS_CHANNEL: "reduce_scatter", ADAM_CHANNEL: "adam", STATS_CHANNEL: "stats"}
        pad = {}
        for ch, nm in names.items():
            for ph in range(3):
                base = (ch * 3 + ph) * int(self.L.bz_comm_max_world())
                pad[f"{nm}.phase{ph}"] = [v & 0xFFFFFFFF for v in tab[base:base + ws]]
        return {"rank": self.e.rank, "host_epochs": {names[k]: v for k, v in self.epoch.items()}, "announced_to_me": pad,
                "k1_launches": self.launches}

    def prepare_debug_snapshot(self) -> None:
        """Create the stream and the pinned buffer ``debug_snapshot`` uses. Call it while the device is healthy: pinning
        host memory can block behind a kernel that never finishes, which is exactly when the snapshot is wanted."""
        if not hasattr(self, "_dbg_stream"):
            self._dbg_stream = torch.cuda.Stream(device=self.e.device)
            self._dbg_host = torch.zeros(self.L.bz_comm_signal_pad_bytes() // 4, dtype=torch.int32).pin_memory()

    def can_push_root(self) -> bool:
        import os
        e = self.e
        return bool(e.stage >= 3 and e.pooled_params and "pbuf_root" in e._symm
                    and os.environ.get("B200ZERO_PUSH_ROOT", "1") == "1")

    def close(self) -> None:
        self.signal.close()
        for b in self.e._symm.values():
            b.close()
        self.e._symm.clear()
