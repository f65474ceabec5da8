"""Checkpoint save / load / verify / rollback in the DeepSpeed on-disk layout.

The reference has no checkpoint code at all — "checkpoint" appears only in remediation strings
and a stub that prints (ref ``ai_engine/spot_resiliency.py:43-49``). The layout
below is the one its README promises to roll back to (``README.md:14``) and the one a DeepSpeed
user expects on disk::

    <dir>/latest                                              text: the newest complete tag
    <dir>/<tag>/mp_rank_00_model_states.pt                    stage ≤ 2 (rank 0): module weights + meta
    <dir>/<tag>/zero_pp_rank_{r}_mp_rank_00_model_states.pt   stage 3: one per rank
    <dir>/<tag>/[bf16_]zero_pp_rank_{r}_mp_rank_00_optim_states.pt   per rank: fp32 master + Adam moments
    <dir>/<tag>/manifest.json                                 sizes + SHA-256 of every file, partition layout
    <dir>/zero_to_fp32.py                                     consolidation helper (ours; understands the layout)

Keys follow DeepSpeed's (``module, optimizer, lr_scheduler, global_steps, global_samples,
skipped_steps, dp_world_size, mp_world_size, ds_config, ds_version, param_shapes`` /
``optimizer_state_dict{zero_stage, partition_count, loss_scaler, dynamic_loss_scale, overflow,
base_optimizer_state, single_partition_of_fp32_groups | fp32_flat_groups}``); the row-partition
layout that differs from DeepSpeed's flat slicing is recorded under ``b200zero_layout`` so any
world size can re-partition on load (elastic resume).

Writes are crash-safe and asynchronous: shards are copied device→pinned host on a side stream,
a background thread serialises them to ``*.tmp``, fsyncs, renames, then drops a per-rank
``.done`` marker; rank 0's writer waits for all markers, writes the manifest and only then
moves ``latest``. A tag without a verifying manifest is never selected for rollback.
"""
from __future__ import annotations

import hashlib
import json
import os
import re
import shutil
import threading
import time
from pathlib import Path
from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

DS_VERSION = "0.13.1+b200zero"
_TAG_RE = re.compile(r"(?:global_step|preempt_step)(\d+)$")


def _dist_ready() -> bool:
    return dist.is_available() and dist.is_initialized()


def _sha256(path: Path) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 22), b""):
            h.update(chunk)
    return h.hexdigest()


def _atomic_torch_save(obj: Any, path: Path) -> None:
    tmp = path.with_name(path.name + f".tmp{os.getpid()}")
    with open(tmp, "wb") as f:
        torch.save(obj, f)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)


def _atomic_write_text(text: str, path: Path) -> None:
    tmp = path.with_name(path.name + f".tmp{os.getpid()}")
    with open(tmp, "w") as f:
        f.write(text)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)


def optim_file(rank: int, bf16: bool) -> str:
    return f"{'bf16_' if bf16 else ''}zero_pp_rank_{rank}_mp_rank_00_optim_states.pt"


def model_file(rank: int, stage: int) -> str:
    return f"zero_pp_rank_{rank}_mp_rank_00_model_states.pt" if stage >= 3 else "mp_rank_00_model_states.pt"


class AsyncWriter:
    """One background thread per process; at most one checkpoint in flight (a second save waits)."""

    def __init__(self):
        self._thread: Optional[threading.Thread] = None
        self.last_error: Optional[BaseException] = None
        self.last_seconds: float = 0.0

    def submit(self, fn) -> None:
        self.wait()

        def run():
            t0 = time.time()
            try:
                fn()
            except BaseException as e:  # surfaced by wait()
                self.last_error = e
            self.last_seconds = time.time() - t0

        self._thread = threading.Thread(target=run, name="ckpt-writer", daemon=False)
        self._thread.start()

    def wait(self) -> None:
        if self._thread is not None:
            self._thread.join()
            self._thread = None
        if self.last_error is not None:
            err, self.last_error = self.last_error, None
            raise RuntimeError(f"asynchronous checkpoint write failed: {err}") from err


_writer = AsyncWriter()


def wait_for_pending_writes() -> None:
    _writer.wait()


def _stage_to_host(engine, tensors: Dict[str, torch.Tensor]) -> Tuple[Dict[str, torch.Tensor], Any]:
    """Copy device tensors into pinned host staging on a side stream (H3); returns
    the host tensors and an event to wait on before serialising."""
    if not engine.on_cuda:
        return {k: v.detach().clone() for k, v in tensors.items()}, None
    if not hasattr(engine, "_ckpt_stream"):
        engine._ckpt_stream = torch.cuda.Stream(device=engine.device)
        engine._ckpt_pinned = {}
    side = engine._ckpt_stream
    side.wait_stream(torch.cuda.current_stream())
    out = {}
    with torch.cuda.stream(side):
        for k, v in tensors.items():
            if not v.is_cuda:
                out[k] = v.detach().clone()
                continue
            key = (k, tuple(v.shape), v.dtype)
            buf = engine._ckpt_pinned.get(key)
            if buf is None:
                buf = torch.empty(v.shape, dtype=v.dtype, pin_memory=True)
                engine._ckpt_pinned[key] = buf
            buf.copy_(v, non_blocking=True)
            out[k] = buf
    evt = torch.cuda.Event()
    evt.record(side)
    return out, evt


def save_checkpoint(engine, save_dir: str, tag: Optional[str] = None, client_state: Optional[Dict] = None,
                    async_write: bool = False, keep_last: Optional[int] = None) -> str:
    cfg = engine.cfg
    # The pinned staging buffers are shared by every save of this engine: a write still in flight (rank 0's writer
    # can sit for minutes waiting for the other ranks' markers) must finish before they are overwritten, otherwise
    # the earlier checkpoint is torn and its checksum — computed after the write — would still verify.
    _writer.wait()
    tag = tag or f"global_step{engine.global_steps}"
    root = Path(save_dir)
    tdir = root / tag
    rank, ws = engine.rank, engine.world_size
    if rank == 0:
        tdir.mkdir(parents=True, exist_ok=True)
        for stale in tdir.glob(".done_rank*"):
            stale.unlink()
        (tdir / "manifest.json").unlink(missing_ok=True)
    if _dist_ready() and ws > 1:
        dist.barrier()
    tdir.mkdir(parents=True, exist_ok=True)
    bf16 = cfg.precision == "bf16"
    layout_desc = engine.layout.describe()

    # --- gather what must be gathered while we are still in lock-step ---
    module_sd = None
    if engine.stage < 3:
        if rank == 0:
            module_sd = engine.module_state_dict()
    elif cfg.gather_16bit_on_save:
        module_sd = engine.module_state_dict()      # collective: every rank participates
        if rank != 0:
            module_sd = None
    staged, evt = _stage_to_host(engine, {"master": engine.master, "exp_avg": engine.exp_avg,
                                          "exp_avg_sq": engine.exp_avg_sq})
    meta_common = {"global_steps": engine.global_steps, "global_samples": engine.global_samples,
                   "skipped_steps": engine.skipped_steps, "dp_world_size": ws, "mp_world_size": 1,
                   "ds_config": cfg.raw, "ds_version": DS_VERSION,
                   "lr_scheduler": engine.schedule.state_dict(), "b200zero_layout": layout_desc}
    param_shapes = [{s["name"]: tuple(s["shape"]) for u in layout_desc["units"] for s in u["params"]}]
    monitor_sd = (client_state or {}).get("monitor")
    step_no = engine.global_steps - engine.skipped_steps
    global_steps = engine.global_steps
    scaler_sd, scaler_dynamic = dict(engine.scaler.state_dict()), engine.scaler.dynamic   # call-time snapshot
    client_state = dict(client_state or {})

    def write() -> None:
        if evt is not None:
            evt.synchronize()
        files: List[Path] = []
        optim_sd = {
            "optimizer_state_dict": {
                "zero_stage": engine.stage, "partition_count": ws, "loss_scaler": scaler_sd,
                "dynamic_loss_scale": scaler_dynamic, "overflow": False,
                "base_optimizer_state": {
                    "state": {0: {"step": step_no, "exp_avg": staged["exp_avg"], "exp_avg_sq": staged["exp_avg_sq"]}},
                    "param_groups": [{"lr": cfg.lr, "betas": tuple(cfg.betas), "eps": cfg.eps,
                                      "weight_decay": cfg.weight_decay, "params": [0]}]},
                ("fp32_flat_groups" if engine.stage >= 3 else "single_partition_of_fp32_groups"): [staged["master"]],
            },
            "ds_config": cfg.raw, "ds_version": DS_VERSION, "param_shapes": param_shapes,
            "b200zero_layout": layout_desc,
        }
        p = tdir / optim_file(rank, bf16)
        _atomic_torch_save(optim_sd, p)
        files.append(p)
        if engine.stage >= 3 or rank == 0:
            msd = dict(meta_common)
            msd.update({"module": module_sd if (rank == 0) else None, "buffer_names": [], "optimizer": None,
                        "param_shapes": param_shapes, "frozen_param_shapes": None, "shared_params": {},
                        "client_state": {k: v for k, v in (client_state or {}).items() if k != "monitor"},
                        "monitor": monitor_sd})
            p = tdir / model_file(rank, engine.stage)
            _atomic_torch_save(msd, p)
            files.append(p)
        entries = {f.name: {"bytes": f.stat().st_size, "sha256": _sha256(f)} for f in files}
        _atomic_write_text(json.dumps(entries), tdir / f".done_rank{rank}")
        if rank == 0:
            deadline = time.time() + 600
            while time.time() < deadline:
                done = [tdir / f".done_rank{r}" for r in range(ws)]
                if all(d.exists() for d in done):
                    break
                time.sleep(0.02)
            else:
                raise TimeoutError(f"ranks did not finish writing {tdir}")
            manifest = {"tag": tag, "global_steps": global_steps, "world_size": ws, "zero_stage": engine.stage,
                        "precision": cfg.precision, "created": time.time(), "files": {}}
            for r in range(ws):
                manifest["files"].update(json.loads((tdir / f".done_rank{r}").read_text()))
            _atomic_write_text(json.dumps(manifest, indent=1), tdir / "manifest.json")
            _atomic_write_text(tag, root / "latest")          # written last: marks the tag complete
            helper = Path(__file__).with_name("zero_to_fp32.py")
            if helper.exists() and not (root / "zero_to_fp32.py").exists():
                shutil.copy(helper, root / "zero_to_fp32.py")
            if keep_last:
                prune_checkpoints(save_dir, keep_last)

    if async_write:
        _writer.submit(write)
    else:
        write()
        if _dist_ready() and ws > 1:
            dist.barrier()
    return str(tdir)


# ---------------------------------------------------------------------- verification / selection
def verify_checkpoint(save_dir: str, tag: str, deep: bool = True) -> Tuple[bool, str]:
    """A tag is good iff its manifest exists and every listed file has the recorded size (and
    SHA-256 when ``deep``) — this is how corrupt checkpoints are identified (README.md:14)."""
    tdir = Path(save_dir) / tag
    mpath = tdir / "manifest.json"
    if not mpath.exists():
        return False, "no manifest (incomplete write)"
    try:
        manifest = json.loads(mpath.read_text())
    except ValueError as e:
        return False, f"manifest unreadable: {e}"
    for name, info in manifest.get("files", {}).items():
        f = tdir / name
        if not f.exists():
            return False, f"missing {name}"
        if f.stat().st_size != info["bytes"]:
            return False, f"size mismatch in {name}"
        if deep and _sha256(f) != info["sha256"]:
            return False, f"checksum mismatch in {name}"
    return True, "ok"


def list_tags(save_dir: str) -> List[Tuple[int, str]]:
    root = Path(save_dir)
    out = []
    if root.exists():
        for d in root.iterdir():
            m = _TAG_RE.search(d.name)
            if d.is_dir() and m:
                out.append((int(m.group(1)), d.name))
    return sorted(out)


def find_rollback_tag(save_dir: str, before_step: Optional[int] = None, deep: bool = True) -> Optional[str]:
    """Newest verified tag with ``global_steps < before_step`` (any step if None)."""
    for step, tag in reversed(list_tags(save_dir)):
        if before_step is not None and step >= before_step:
            continue
        ok, _ = verify_checkpoint(save_dir, tag, deep=deep)
        if ok:
            return tag
    return None


def prune_checkpoints(save_dir: str, keep_last: int) -> List[str]:
    tags = list_tags(save_dir)
    removed = []
    for _, tag in tags[:-keep_last] if keep_last > 0 else []:
        shutil.rmtree(Path(save_dir) / tag, ignore_errors=True)
        removed.append(tag)
    return removed


# ---------------------------------------------------------------------- load (with re-partitioning)
def _load_optim_files(tdir: Path, old_ws: int, bf16: bool) -> List[Dict[str, Any]]:
    out = []
    for r in range(old_ws):
        p = tdir / optim_file(r, bf16)
        if not p.exists():
            p = tdir / optim_file(r, not bf16)
        out.append(torch.load(p, map_location="cpu", weights_only=False))
    return out


def consolidate_fp32(optim_sds: List[Dict[str, Any]], which: str = "master") -> Dict[str, torch.Tensor]:
    """Full fp32 tensors by name from all ranks' shard files (``which``: master|exp_avg|exp_avg_sq)."""
    layout = optim_sds[0]["b200zero_layout"]
    ws = layout["world_size"]
    flats = []
    for sd in optim_sds:
        osd = sd["optimizer_state_dict"]
        if which == "master":
            flats.append((osd.get("fp32_flat_groups") or osd.get("single_partition_of_fp32_groups"))[0])
        else:
            flats.append(osd["base_optimizer_state"]["state"][0][which])
    full: Dict[str, torch.Tensor] = {}
    for u in layout["units"]:
        for s in u["params"]:
            off, n, shape = s["shard_offset"], s["shard_numel"], tuple(s["shape"])
            if s["partitioned"]:
                rows = [flats[r][off:off + n].view((shape[0] // ws,) + shape[1:]) for r in range(ws)]
                full[s["name"]] = torch.cat(rows, 0).clone()
            else:
                full[s["name"]] = flats[0][off:off + n].view(shape).clone()
    return full


def load_checkpoint(engine, load_dir: str, tag: Optional[str] = None, verify: bool = True) -> Tuple[Optional[str], Dict]:
    root = Path(load_dir)
    if tag is None:
        latest = root / "latest"
        if not latest.exists():
            return None, {}
        tag = latest.read_text().strip()
    if verify:
        ok, why = verify_checkpoint(load_dir, tag, deep=True)
        if not ok:
            raise RuntimeError(f"checkpoint {tag} failed verification: {why}")
    tdir = root / tag
    manifest = json.loads((tdir / "manifest.json").read_text())
    old_ws, old_stage = manifest["world_size"], manifest["zero_stage"]
    bf16 = manifest.get("precision") == "bf16"
    mpath = tdir / model_file(0, old_stage)
    msd = torch.load(mpath, map_location="cpu", weights_only=False)
    same_layout = (old_ws == engine.world_size and
                   msd["b200zero_layout"]["shard_numel"] == engine.layout.shard_numel and
                   msd["b200zero_layout"]["part_numel"] == engine.layout.part_numel)
    if same_layout:
        sd = torch.load(tdir / optim_file(engine.rank, bf16), map_location="cpu", weights_only=False)
        osd = sd["optimizer_state_dict"]
        master = (osd.get("fp32_flat_groups") or osd.get("single_partition_of_fp32_groups"))[0]
        st = osd["base_optimizer_state"]["state"][0]
        engine.master.copy_(master)
        engine.exp_avg.copy_(st["exp_avg"])
        engine.exp_avg_sq.copy_(st["exp_avg_sq"])
        engine.scaler.load_state_dict(osd.get("loss_scaler", {}))
    else:
        # elastic resume: rebuild full fp32 tensors from the old partitioning, re-slice for ours
        sds = _load_optim_files(tdir, old_ws, bf16)
        for which, dst in (("master", engine.master), ("exp_avg", engine.exp_avg), ("exp_avg_sq", engine.exp_avg_sq)):
            full = consolidate_fp32(sds, which)
            for u in engine.layout.units:
                for spec in u.params:
                    src = engine._slice_rows(spec, full[spec.name])
                    dst[spec.shard_offset:spec.shard_offset + spec.shard_numel].copy_(src)
        engine.scaler.load_state_dict(sds[0]["optimizer_state_dict"].get("loss_scaler", {}))
    engine.global_steps = int(msd["global_steps"])
    engine.global_samples = int(msd.get("global_samples", 0))
    engine.skipped_steps = int(msd.get("skipped_steps", 0))
    engine.micro_steps = engine.global_steps * engine.cfg.grad_accum
    engine.schedule.load_state_dict(msd.get("lr_scheduler", {}))
    engine.schedule.last_step = engine.global_steps
    engine._refresh_shard_params()
    engine.shard_grads.zero_()
    for gb in {b.data_ptr(): b for b in engine._unit_gbuf}.values():
        gb.zero_()
    engine._loss_accum.zero_()
    if _dist_ready() and engine.world_size > 1:
        # the fused gathers read the peers' shard_params directly over NVLink: every rank's restored rows must be
        # in memory before anyone gathers (same rule as ZeroEngine.load_full_params)
        if engine.on_cuda:
            torch.cuda.synchronize(engine.device)
        dist.barrier()
    if engine.stage >= 3 and engine.pooled_params:
        engine._resident = [False] * len(engine.layout.units)
        engine._slot_owner = [None] * engine._n_pslots
        engine._gather_unit(0)
    else:
        for ui in range(len(engine.layout.units)):
            engine._gather_unit(ui)
    client = dict(msd.get("client_state") or {})
    if msd.get("monitor") is not None:
        client["monitor"] = msd["monitor"]
    if _dist_ready() and engine.world_size > 1:
        dist.barrier()
    return str(tdir), client
