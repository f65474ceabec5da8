"""Flag/env-driven fault injection. The reference's only injection
mechanism is editing ``spot_resiliency.py:41`` by hand. Every fault fires at most once per
process lifetime *per global step*, so a rolled-back run that replays the step is clean —
which is what a transient hardware fault looks like.

    B200ZERO_INJECT_SPIKE_AT=<step>[:<factor>]   multiply the loss (and so every gradient)
    B200ZERO_INJECT_NAN_GRAD_AT=<step>           poison one gradient element with NaN
    B200ZERO_KILL_RANK_AT=<step>:<rank>          SIGKILL that rank at the top of the step
    B200ZERO_PREEMPT_AT=<step>                   deliver SIGTERM to self (spot preemption)
    B200ZERO_HANG_AT=<step>:<rank>               that rank stops at the top of the step and never returns (a rank
                                                 stuck inside a collective); first start only, like the kill
"""
from __future__ import annotations

import os
import signal
from dataclasses import dataclass, field
from typing import Callable, Optional, Set


@dataclass
class FaultPlan:
    spike_at: Optional[int] = None
    spike_factor: float = 50.0
    nan_grad_at: Optional[int] = None
    nan_grad_unit: Optional[int] = None     # poison that unit's gradient bucket just before it is reduced (mid-backward)
    kill_at: Optional[int] = None
    kill_rank: int = 0
    preempt_at: Optional[int] = None
    hang_at: Optional[int] = None
    hang_rank: int = 0
    on_inject: Optional[Callable] = None    # called right after a mid-backward fault was written (probes timestamp it)
    fired: Set[str] = field(default_factory=set)

    @classmethod
    def from_env(cls, env=os.environ) -> "FaultPlan":
        p = cls()
        v = env.get("B200ZERO_INJECT_SPIKE_AT")
        if v:
            a, _, b = v.partition(":")
            p.spike_at, p.spike_factor = int(a), float(b) if b else 50.0
        v = env.get("B200ZERO_INJECT_NAN_GRAD_AT")
        if v:
            p.nan_grad_at = int(v)
        v = env.get("B200ZERO_KILL_RANK_AT")
        if v:
            a, _, b = v.partition(":")
            p.kill_at, p.kill_rank = int(a), int(b or 0)
        v = env.get("B200ZERO_PREEMPT_AT")
        if v:
            p.preempt_at = int(v)
        v = env.get("B200ZERO_HANG_AT")
        if v:
            a, _, b = v.partition(":")
            p.hang_at, p.hang_rank = int(a), int(b or 0)
        return p

    def _once(self, key: str) -> bool:
        if key in self.fired:
            return False
        self.fired.add(key)
        return True

    def install(self, engine) -> None:
        plan = self

        def pre_backward(eng, loss):
            step = eng.global_steps + 1
            if plan.spike_at == step and plan._once(f"spike{step}"):
                return loss * plan.spike_factor
            return loss

        def pre_reduce(eng, ui):
            # a fault that appears while the backward pass is still running: the NaN sits in the unit's gradient
            # bucket when its reduce-scatter starts, so the reduce kernels (or the norm pass) flag it themselves
            step = eng.global_steps + 1
            if plan.nan_grad_at == step and plan.nan_grad_unit == ui and plan._once(f"nan{step}"):
                eng._unit_gbuf[ui][:1] = float("nan")
                if plan.on_inject is not None:
                    plan.on_inject(eng)

        def post_backward(eng):
            step = eng.global_steps + 1
            if plan.nan_grad_at == step and plan.nan_grad_unit is None and plan._once(f"nan{step}"):
                eng.shard_grads[0] = float("nan")
                if getattr(eng, "fused", None) is not None:   # the fused path derives the flag inside its reduce kernels
                    eng._flag.fill_(1)

        engine.fault_hooks["pre_backward"] = pre_backward
        engine.fault_hooks["post_backward"] = post_backward
        engine.fault_hooks["pre_reduce"] = pre_reduce

    def uninstall(self, engine) -> None:
        for k in ("pre_backward", "post_backward", "pre_reduce"):
            engine.fault_hooks.pop(k, None)

    def at_step_start(self, engine) -> None:
        step = engine.global_steps + 1
        if self.kill_at == step and engine.rank == self.kill_rank and self._once(f"kill{step}") \
                and os.environ.get("B200ZERO_RESTART_COUNT", "0") == "0":
            os.kill(os.getpid(), signal.SIGKILL)
        if self.hang_at == step and engine.rank == self.hang_rank and self._once(f"hang{step}") \
                and os.environ.get("B200ZERO_RESTART_COUNT", "0") == "0":
            import time
            while True:                      # the other ranks now wait for this one inside their next collective
                time.sleep(3600)
        if self.preempt_at == step and self._once(f"preempt{step}") \
                and os.environ.get("B200ZERO_RESTART_COUNT", "0") == "0":
            os.kill(os.getpid(), signal.SIGTERM)
