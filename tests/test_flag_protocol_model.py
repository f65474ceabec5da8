"""CPU model of the cross-GPU flag protocol used by the fused collectives (csrc/fused_comm.cu), with
threads as ranks. It checks the *logic* of the protocol — epochs, the two barrier
phases of the reduce-scatter kernel (K1), the gradient-slot reuse rule, and the post-Adam barrier of the
push all-gather (K3) — under adversarial random delays; memory-model details (st.release.sys /
ld.acquire.sys, __threadfence_system) are the kernel's business and are covered by the multi-GPU tests.

Model:  each rank has a "main stream" thread (backward, optimizer, forward) and a "comm stream" thread
(K1 launches, in order). Buffers carry tags instead of numbers, so every read can assert that it sees
exactly the (step, unit) it is supposed to see, fully written.
"""
import queue
import random
import threading
import time

import pytest

UNITS, STEPS = 5, 4
RS, ADAM, STATS = 0, 1, 2


class World:
    def __init__(self, ws, k1_phase1=True, k3_phase2=True, seed=0, stats_bufs=2):
        self.ws = ws
        self.k1_phase1, self.k3_phase2, self.stats_bufs = k1_phase1, k3_phase2, stats_bufs
        # payload[r][buf][src]: the control vector src published to r for one exchange
        self.payload = [[[None] * ws for _ in range(2)] for _ in range(ws)]
        # signal[r][(channel, phase)][src] = last epoch announced by src to r
        self.signal = [{(c, p): [0] * ws for c in (RS, ADAM, STATS) for p in (0, 1)} for _ in range(ws)]
        self.grad_slot = [[None, None] for _ in range(ws)]            # two pooled gradient buffers per rank
        self.params = [[None] * ws for _ in range(ws)]                # params[r][owner] = tag of owner's rows as seen by r
        self.errors = []
        self.rng = [random.Random(seed * 100 + r) for r in range(ws)]
        self.deadline = time.time() + 60

    def nap(self, r, scale=1.0):
        time.sleep(self.rng[r].random() * 0.002 * scale)

    def signal_all_and_wait(self, r, channel, phase, epoch):
        for dst in range(self.ws):                                    # st.release.sys to every peer's pad
            self.signal[dst][(channel, phase)][r] = epoch
        mine = self.signal[r][(channel, phase)]
        while any(mine[src] - epoch < 0 for src in range(self.ws)):   # ld.acquire.sys, epochs only grow
            if time.time() > self.deadline:
                raise TimeoutError(f"rank {r} stuck on channel {channel} phase {phase} epoch {epoch}: {mine}")
            time.sleep(0)

    def fail(self, msg):
        self.errors.append(msg)


def k1(w, r, step, unit, epoch):
    """Reduce-scatter of `unit`: pull my rows from every peer's slot."""
    slot = unit % 2
    w.signal_all_and_wait(r, RS, 0, epoch)                            # phase 0: everyone's gradients are ready
    for p in range(w.ws):
        w.nap(r, 0.5)
        tag = w.grad_slot[p][slot]
        if tag != (step, unit, "complete"):
            w.fail(f"rank {r} read {tag} from rank {p} slot {slot}, wanted {(step, unit)}")
    if w.k1_phase1:
        w.signal_all_and_wait(r, RS, 1, epoch)                        # phase 1: everyone finished reading my slot


def k3(w, r, step, epoch):
    """Adam on my shard, push my rows into every rank's parameter buffer, then barrier."""
    for dst in range(w.ws):
        w.nap(r, 0.3)
        w.params[dst][r] = step
    if w.k3_phase2:
        w.signal_all_and_wait(r, ADAM, 1, epoch)


def stats_exchange(w, r, step, epoch):
    """bz_peer_allreduce4: payload into every peer's pad, flag, wait for all flags, read all payloads."""
    buf = epoch % w.stats_bufs
    for dst in range(w.ws):
        w.payload[dst][buf][r] = step
        w.nap(r, 0.2)
        w.signal[dst][(STATS, 0)][r] = epoch
    mine = w.signal[r][(STATS, 0)]
    while any(mine[src] - epoch < 0 for src in range(w.ws)):
        if time.time() > w.deadline:
            raise TimeoutError(f"rank {r} stuck in the stats exchange, epoch {epoch}: {mine}")
        time.sleep(0)
    w.nap(r, 2.0)                                                    # adversarial: slow reader
    for src in range(w.ws):
        if w.payload[r][buf][src] != step:
            w.fail(f"rank {r} summed rank {src}'s vector of step {w.payload[r][buf][src]} in step {step}")


def rank_main(w, r):
    comm_q = queue.Queue()
    slot_free = [threading.Event(), threading.Event()]
    for e in slot_free:
        e.set()
    comm_done = threading.Event()

    def comm_stream():
        try:
            while True:
                item = comm_q.get()
                if item is None:
                    return
                step, unit, epoch = item
                k1(w, r, step, unit, epoch)
                slot_free[unit % 2].set()                             # kernel completion = slot may be overwritten
                if unit == 0:
                    comm_done.set()
        except Exception as exc:  # noqa: BLE001
            w.fail(f"rank {r} comm stream: {exc!r}")
            comm_done.set()
            for e in slot_free:
                e.set()

    t = threading.Thread(target=comm_stream, daemon=True)
    t.start()
    rs_epoch = adam_epoch = 0
    try:
        for step in range(1, STEPS + 1):
            # forward: my parameter buffer must hold every owner's rows of the previous step
            if step > 1:
                for owner in range(w.ws):
                    if w.params[r][owner] != step - 1:
                        w.fail(f"rank {r} forward of step {step} saw rows of rank {owner} at {w.params[r][owner]}")
            w.nap(r, 3.0)
            comm_done.clear()
            for unit in range(UNITS - 1, -1, -1):                     # backward, last unit first
                slot = unit % 2
                slot_free[slot].wait()                                # wait_event(_gslot_free_evt[slot])
                slot_free[slot].clear()
                w.grad_slot[r][slot] = (step, unit, "partial")        # backward kernels write the slot ...
                w.nap(r, 2.0)
                w.grad_slot[r][slot] = (step, unit, "complete")
                rs_epoch += 1
                comm_q.put((step, unit, rs_epoch))                    # ... then K1 is enqueued behind them
            comm_done.wait()                                          # main stream waits for the comm stream
            stats_exchange(w, r, step, step)
            adam_epoch += 1
            k3(w, r, step, adam_epoch)
    except Exception as exc:  # noqa: BLE001
        w.fail(f"rank {r} main stream: {exc!r}")
    finally:
        comm_q.put(None)
        t.join(10)


def run_world(ws, **kw):
    w = World(ws, **kw)
    threads = [threading.Thread(target=rank_main, args=(w, r), daemon=True) for r in range(ws)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(90)
    assert not any(t.is_alive() for t in threads), "model deadlocked"
    return w


@pytest.mark.parametrize("ws,seed", [(2, 0), (4, 1), (8, 2)])
def test_protocol_is_safe_and_live_under_random_delays(ws, seed):
    w = run_world(ws, seed=seed)
    assert w.errors == []


def test_model_has_teeth_without_k1_phase1():
    """Without the 'everyone finished reading' phase a fast rank recycles a gradient slot under a slow
    reader: the model must catch that, otherwise it proves nothing."""
    caught = False
    for seed in range(6):
        w = run_world(4, k1_phase1=False, seed=seed)
        if any("read" in e for e in w.errors):
            caught = True
            break
    assert caught


def test_model_has_teeth_without_k3_barrier():
    caught = False
    for seed in range(6):
        w = run_world(4, k3_phase2=False, seed=seed)
        if any("forward" in e for e in w.errors):
            caught = True
            break
    assert caught


def run_exchange_only(ws, stats_bufs, seed, steps=8):
    """Back-to-back bz_peer_allreduce4 calls with nothing else between them (the kernel's own contract, without
    the K3 barrier that happens to follow it in the engine)."""
    w = World(ws, seed=seed, stats_bufs=stats_bufs)

    def rank(r):
        try:
            for step in range(1, steps + 1):
                stats_exchange(w, r, step, step)
        except Exception as exc:  # noqa: BLE001
            w.fail(f"rank {r}: {exc!r}")
    threads = [threading.Thread(target=rank, args=(r,), daemon=True) for r in range(ws)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(60)
    assert not any(t.is_alive() for t in threads), "model deadlocked"
    return w


@pytest.mark.parametrize("ws", [2, 4, 8])
def test_stats_exchange_is_safe_back_to_back(ws):
    for seed in range(3):
        assert run_exchange_only(ws, 2, seed).errors == []


def test_model_has_teeth_without_stats_double_buffer():
    """One payload buffer is not enough: a fast rank's next vector lands before a slow rank has read this one."""
    caught = False
    for seed in range(6):
        w = run_exchange_only(4, 1, seed)
        if any("summed" in e for e in w.errors):
            caught = True
            break
    assert caught
