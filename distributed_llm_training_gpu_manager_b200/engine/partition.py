"""ZeRO partition layout.

Every parameter is either **row-partitioned** (dim 0 split evenly over the data-parallel ranks;
rank r owns rows ``[r·n/ws, (r+1)·n/ws)``) or **replicated** (small tensors below
``stage3_param_persistence_threshold`` elements, or whose dim 0 does not divide — norm gains,
biases; ref ``deepspeed_launcher.py:191`` emits that threshold). Row partitioning — rather than
DeepSpeed's contiguous slicing of the flat concatenation — is what lets the fused kernels work
tile by tile: a gathered weight is contiguous in its natural [N, K] layout, every GEMM N-tile has
exactly one owner, traffic to any one matrix is spread evenly over all peers, and the gradient
shard has the same layout as the parameter shard so the optimizer is purely element-wise.

Parameters are grouped into **units** (one transformer block each, plus a root unit) — the
granularity of all-gather / reduce-scatter and of ``stage3_max_live_parameters`` accounting.

Per-rank flat buffers all share one *shard layout*:
    [ partitioned slices, unit by unit | replicated tensors, unit by unit ]
and each unit has a *full layout* (its gathered parameters back to back). All offsets are
multiples of 16 elements so every tensor base is TMA/vector aligned in any dtype.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Sequence, Tuple

ALIGN = 16


def _up(n: int, a: int = ALIGN) -> int:
    return (n + a - 1) // a * a


@dataclass
class ParamSpec:
    name: str
    shape: Tuple[int, ...]
    unit: int
    partitioned: bool = False
    full_offset: int = 0     # within the unit's full buffer
    shard_offset: int = 0    # within the rank's shard buffer
    shard_numel: int = 0     # elements this rank holds (slice or whole tensor)

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n

    def shard_shape(self, ws: int) -> Tuple[int, ...]:
        if not self.partitioned:
            return self.shape
        return (self.shape[0] // ws,) + tuple(self.shape[1:])


@dataclass
class UnitLayout:
    index: int
    name: str
    params: List[ParamSpec] = field(default_factory=list)
    full_numel: int = 0
    part_shard_start: int = 0   # [start, end) of this unit's partitioned slices in the shard buffer
    part_shard_end: int = 0
    repl_shard_start: int = 0
    repl_shard_end: int = 0


@dataclass
class Layout:
    world_size: int
    units: List[UnitLayout]
    shard_numel: int          # total elements per rank
    part_numel: int           # [0, part_numel) = partitioned region, rest = replicated region
    by_name: Dict[str, ParamSpec]

    @property
    def total_params(self) -> int:
        return sum(p.numel for u in self.units for p in u.params)

    def max_unit_numel(self) -> int:
        return max((u.full_numel for u in self.units), default=0)

    def describe(self) -> Dict:
        return {"world_size": self.world_size, "shard_numel": self.shard_numel, "part_numel": self.part_numel,
                "units": [{"name": u.name, "full_numel": u.full_numel,
                           "params": [{"name": p.name, "shape": list(p.shape), "partitioned": p.partitioned,
                                       "full_offset": p.full_offset, "shard_offset": p.shard_offset,
                                       "shard_numel": p.shard_numel} for p in u.params]} for u in self.units]}


def build_layout(params: Sequence[Tuple[str, Tuple[int, ...], int]], unit_names: Sequence[str], world_size: int,
                 partition: bool, persistence_threshold: int = 100_000) -> Layout:
    """``params``: (name, shape, unit_index) in execution order. ``partition=False`` (ZeRO-0)
    replicates everything."""
    units = [UnitLayout(i, n) for i, n in enumerate(unit_names)]
    by_name: Dict[str, ParamSpec] = {}
    for name, shape, ui in params:
        numel = 1
        for s in shape:
            numel *= s
        part = (partition and world_size > 1 and len(shape) >= 1 and shape[0] % world_size == 0
                and numel >= persistence_threshold and (numel // world_size) % ALIGN == 0)
        spec = ParamSpec(name, tuple(shape), ui, part)
        units[ui].params.append(spec)
        by_name[name] = spec
    # full layouts
    for u in units:
        off = 0
        for p in u.params:
            p.full_offset = off
            off += _up(p.numel)
        u.full_numel = off
    # shard layout: partitioned region first, then replicated
    off = 0
    for u in units:
        u.part_shard_start = off
        for p in u.params:
            if p.partitioned:
                p.shard_offset = off
                p.shard_numel = p.numel // world_size
                off += _up(p.shard_numel)
        u.part_shard_end = off
    part_numel = off
    for u in units:
        u.repl_shard_start = off
        for p in u.params:
            if not p.partitioned:
                p.shard_offset = off
                p.shard_numel = p.numel
                off += _up(p.numel)
        u.repl_shard_end = off
    return Layout(world_size, units, off, part_numel, by_name)
