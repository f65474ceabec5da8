"""Learning-rate schedules with DeepSpeed's names and parameter keys (the launcher emits
``WarmupDecayLR`` with ``warmup_min_lr / warmup_max_lr / warmup_num_steps / total_num_steps``,
ref ``deepspeed_launcher.py:145-153``), fp16 dynamic loss scaling (ref ``:175-183``), and the
elastic batch solver for the ``elasticity`` block (ref ``:226-238``)."""
from __future__ import annotations

import math
from typing import Any, Dict, List, Optional, Tuple


class LRSchedule:
    def __init__(self, kind: Optional[str], params: Dict[str, Any], base_lr: float):
        self.kind = kind or "Constant"
        self.p = dict(params)
        self.base_lr = base_lr
        self.last_step = 0

    def lr_at(self, step: int) -> float:
        """Learning rate for optimizer step ``step`` (0-based count of completed steps)."""
        p = self.p
        lo = float(p.get("warmup_min_lr", 0.0))
        hi = float(p.get("warmup_max_lr", self.base_lr))
        warm = int(p.get("warmup_num_steps", 0))
        kind = self.kind
        if kind == "Constant":
            return self.base_lr
        if step < warm:
            if p.get("warmup_type", "log") == "linear" or kind == "WarmupDecayLR" and p.get("warmup_type") == "linear":
                frac = step / max(warm, 1)
            else:  # DeepSpeed's default warm-up is logarithmic
                frac = math.log(step + 1) / math.log(max(warm, 2))
            return lo + (hi - lo) * min(max(frac, 0.0), 1.0)
        if kind == "WarmupLR":
            return hi
        total = int(p.get("total_num_steps", warm + 1))
        if kind == "WarmupDecayLR":
            return hi * max(0.0, (total - step) / max(total - warm, 1))
        if kind == "WarmupCosineLR":
            ratio = float(p.get("cos_min_ratio", 0.0001))
            prog = min(1.0, (step - warm) / max(total - warm, 1))
            return hi * (ratio + (1 - ratio) * 0.5 * (1 + math.cos(math.pi * prog)))
        raise ValueError(f"unsupported scheduler type {kind!r}")

    def step(self) -> float:
        self.last_step += 1
        return self.lr_at(self.last_step)

    def state_dict(self) -> Dict[str, Any]:
        return {"last_batch_iteration": self.last_step - 1, "kind": self.kind, "params": self.p}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        self.last_step = int(sd.get("last_batch_iteration", -1)) + 1


class DynamicLossScaler:
    """fp16 loss scaling: halve on overflow (after ``hysteresis`` consecutive overflows), double
    after ``window`` clean steps. Static when ``loss_scale > 0``. Inert (scale 1) for bf16/fp32."""

    def __init__(self, enabled: bool, static_scale: float = 0.0, init_power: int = 16, window: int = 1000,
                 hysteresis: int = 2, min_scale: float = 1.0):
        self.enabled = enabled
        self.dynamic = enabled and static_scale == 0
        self.scale = (static_scale if static_scale > 0 else 2.0 ** init_power) if enabled else 1.0
        self.window, self.hysteresis, self.min_scale = window, hysteresis, min_scale
        self._hyst_left = hysteresis
        self._good = 0
        self.skipped = 0

    def update(self, overflow: bool) -> None:
        if overflow:
            self.skipped += 1
        if not self.dynamic:
            return
        if overflow:
            self._good = 0
            if self._hyst_left > 1:
                self._hyst_left -= 1
            else:
                self.scale = max(self.scale / 2.0, self.min_scale)
        else:
            self._good += 1
            if self._good % self.window == 0:
                self.scale *= 2.0
                self._hyst_left = self.hysteresis

    def state_dict(self) -> Dict[str, Any]:
        return {"cur_scale": self.scale, "cur_iter": self._good, "cur_hysteresis": self._hyst_left,
                "skipped": self.skipped, "dynamic": self.dynamic}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        self.scale = float(sd.get("cur_scale", self.scale))
        self._good = int(sd.get("cur_iter", 0))
        self._hyst_left = int(sd.get("cur_hysteresis", self.hysteresis))
        self.skipped = int(sd.get("skipped", 0))


def elastic_batch_plan(elastic: Dict[str, Any], micro_batches: Optional[List[int]] = None) -> Tuple[int, List[int]]:
    """Pick the global batch (≤ ``max_train_batch_size``) compatible with the most GPU counts in
    [min_gpus, max_gpus] and return ``(batch, valid_gpu_counts)``: a world size ``n`` is valid when
    ``batch % (micro · n) == 0`` for some allowed micro batch. Resume at any valid ``n`` keeps the
    global batch (and hence the LR schedule) intact."""
    hi = int(elastic.get("max_train_batch_size", 0))
    lo_mb = int(elastic.get("min_train_batch_size", 1))
    gmin, gmax = int(elastic.get("min_gpus", 1)), int(elastic.get("max_gpus", 1))
    micro = sorted(set(micro_batches or [lo_mb]))
    best: Tuple[int, int, List[int]] = (-1, 0, [])
    for batch in range(hi, 0, -1):
        valid = [n for n in range(gmin, gmax + 1) if any(batch % (mb * n) == 0 for mb in micro)]
        key = (len(valid), batch if elastic.get("prefer_larger_batch_size", True) else -batch)
        if key > (best[0], best[1] if elastic.get("prefer_larger_batch_size", True) else -best[1]):
            best = (len(valid), batch, valid)
        if len(valid) == gmax - gmin + 1:
            break
    return best[1], best[2]
