"""Token-file input pipeline for the packaged worker and the Trainer.

The reference ships no data path at all (its launcher starts a user script it does not include), and
the benchmarks use synthetic tokens; this is what a user needs to train on real data with the native engine:

* ``TokenFileDataset`` — a flat binary file of token ids (uint16 / uint32 / int32, e.g. the output of any
  "tokenize to .bin" script), memory-mapped, cut into non-overlapping windows of ``seq_len + 1`` tokens.
* ``token_batches`` — a ``batches(micro_step) -> (ids, labels)`` callable in the Trainer's protocol. The window
  indices of a micro-step are a pure function of ``(seed, epoch, micro_step, rank)``: every rank reads a disjoint
  slice of a per-epoch permutation, and a resumed or rolled-back run replays exactly the batches it saw before
  (the Trainer asks by ``engine.micro_steps``). A background thread assembles the next micro-steps into pinned
  host buffers and issues their host→device copies on a side stream, so the training stream only waits on an event.
"""
from __future__ import annotations

import threading
from pathlib import Path
from typing import Callable, Dict, Optional, Tuple

import numpy as np
import torch

_DTYPES = {"uint16": np.uint16, "uint32": np.uint32, "int32": np.int32, "int64": np.int64}


class TokenFileDataset:
    def __init__(self, path, seq_len: int, dtype: str = "auto", vocab_size: Optional[int] = None):
        self.path = Path(path)
        self.seq_len = int(seq_len)
        nbytes = self.path.stat().st_size
        if dtype == "auto":
            # the common convention: 2-byte ids when the vocabulary fits, else 4-byte
            dtype = "uint16" if (vocab_size is not None and vocab_size <= 65536) else "uint32"
        if dtype not in _DTYPES:
            raise ValueError(f"unknown token dtype {dtype!r}; expected one of {sorted(_DTYPES)}")
        self.dtype = np.dtype(_DTYPES[dtype])
        if nbytes % self.dtype.itemsize:
            raise ValueError(f"{self.path}: size {nbytes} is not a multiple of {self.dtype.itemsize}-byte tokens")
        self.tokens = np.memmap(self.path, dtype=self.dtype, mode="r")
        self.n_windows = (len(self.tokens) - 1) // self.seq_len
        if self.n_windows < 1:
            raise ValueError(f"{self.path}: {len(self.tokens)} tokens are fewer than one window of {self.seq_len + 1}")
        self.vocab_size = vocab_size

    def __len__(self) -> int:
        return self.n_windows

    def window(self, idx: int) -> np.ndarray:
        """``seq_len + 1`` tokens starting at window ``idx`` (consecutive windows share one boundary token, so
        every token is a target exactly once)."""
        o = idx * self.seq_len
        return self.tokens[o:o + self.seq_len + 1]


def _epoch_permutation(n: int, seed: int, epoch: int) -> np.ndarray:
    return np.random.Generator(np.random.PCG64([seed, epoch])).permutation(n)


class _Plan:
    """Maps a micro-step to the window indices of one rank."""

    def __init__(self, n_windows: int, micro_batch: int, rank: int, world_size: int, seed: int, shuffle: bool):
        self.n, self.mb, self.rank, self.ws, self.seed, self.shuffle = n_windows, micro_batch, rank, world_size, seed, shuffle
        self.per_step = micro_batch * world_size
        self.steps_per_epoch = max(self.n // self.per_step, 1)
        self._perm: Dict[int, np.ndarray] = {}

    def indices(self, micro_step: int) -> np.ndarray:
        epoch, k = divmod(micro_step, self.steps_per_epoch)
        if self.shuffle:
            if epoch not in self._perm:
                self._perm = {epoch: _epoch_permutation(self.n, self.seed, epoch)}
            order = self._perm[epoch]
        else:
            order = None
        base = k * self.per_step + self.rank * self.mb
        pos = (base + np.arange(self.mb)) % self.n
        return order[pos] if order is not None else pos


def token_batches(path, micro_batch: int, seq_len: int, device, rank: int = 0, world_size: int = 1, seed: int = 1234,
                  shuffle: bool = True, dtype: str = "auto", vocab_size: Optional[int] = None,
                  prefetch: int = 2) -> Callable[[int], Tuple[torch.Tensor, torch.Tensor]]:
    ds = TokenFileDataset(path, seq_len, dtype=dtype, vocab_size=vocab_size)
    plan = _Plan(len(ds), micro_batch, rank, world_size, seed, shuffle)
    dev = torch.device(device)
    on_cuda = dev.type == "cuda"
    side = torch.cuda.Stream(device=dev) if on_cuda else None
    lock = threading.Lock()
    ready: Dict[int, Tuple[torch.Tensor, torch.Tensor, Optional[torch.cuda.Event]]] = {}
    wanted = threading.Event()
    state = {"next": 0, "stop": False}

    def assemble(micro_step: int):
        idx = plan.indices(micro_step)
        host = torch.empty((micro_batch, seq_len + 1), dtype=torch.int64, pin_memory=on_cuda)
        hn = host.numpy()
        for b, i in enumerate(idx):
            hn[b] = ds.window(int(i))
        if vocab_size is not None and int(hn.max()) >= vocab_size:
            raise ValueError(f"{ds.path}: token id {int(hn.max())} >= vocab_size {vocab_size}")
        if not on_cuda:
            return host[:, :-1].contiguous(), host[:, 1:].contiguous(), None
        with torch.cuda.stream(side):
            d = host.to(dev, non_blocking=True)
            ids, labels = d[:, :-1].contiguous(), d[:, 1:].contiguous()
            evt = torch.cuda.Event()
            evt.record(side)
        return ids, labels, evt, host          # the pinned buffer stays referenced until the batch is consumed

    def worker():
        if on_cuda:
            torch.cuda.set_device(dev)
        while not state["stop"]:
            with lock:
                todo = [s for s in range(state["next"], state["next"] + prefetch) if s not in ready]
            if not todo:
                wanted.wait(0.05)
                wanted.clear()
                continue
            item = assemble(todo[0])
            with lock:
                ready[todo[0]] = item

    t = threading.Thread(target=worker, daemon=True, name="token-prefetch")
    t.start()

    def get(micro_step: int) -> Tuple[torch.Tensor, torch.Tensor]:
        with lock:
            item = ready.pop(micro_step, None)
            for s in [s for s in ready if s < micro_step]:       # a rollback moved the cursor: drop stale batches
                ready.pop(s)
            state["next"] = micro_step + 1
        wanted.set()
        if item is None:
            item = assemble(micro_step)                          # first call, or a jump the prefetcher had not seen
        ids, labels, evt = item[0], item[1], item[2]
        if evt is not None:
            torch.cuda.current_stream().wait_event(evt)
            ids.record_stream(torch.cuda.current_stream())
            labels.record_stream(torch.cuda.current_stream())
        return ids, labels

    get.dataset = ds                     # type: ignore[attr-defined]
    get.steps_per_epoch = plan.steps_per_epoch   # type: ignore[attr-defined]
    get.close = lambda: state.update(stop=True)   # type: ignore[attr-defined]
    return get
