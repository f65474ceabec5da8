"""CPU model of the mbarrier protocol of the pipelined flash-attention backward kernel (csrc/attention_bwd_v3.cu).

v2 runs the five GEMMs of a (KV block, query tile) pair strictly after one another with the compute warps in
between. v3 issues S^T of tile i+1 right after dV of tile i (double-buffered S^T columns in TMEM, three Q/dO
stages) and lets the compute warps stage dQ^T of tile i-1 *after* dS^T of tile i, so the tensor core and the
compute warps overlap. The price is a protocol with more single-buffered resources guarded by "consumer done"
barriers (dS^T smem by dq_full, the dQ^T columns by dqt_empty; P^T smem and the dP^T columns by the in-order
tensor pipe / program order). This model executes that protocol with hardware mbarrier semantics (phase bit, aliasing two phases
behind), an in-order tensor pipe, tagged buffers that assert on every read/overwrite, and random delays.
"""
import queue
import random
import threading
import time

import pytest


class MBar:
    def __init__(self, count):
        self.count, self.pending, self.phase = count, count, 0
        self.cv = threading.Condition()

    def arrive(self):
        with self.cv:
            self.pending -= 1
            if self.pending == 0:
                self.pending = self.count
                self.phase += 1
                self.cv.notify_all()

    def wait(self, parity, deadline):
        with self.cv:
            while (self.phase & 1) == (parity & 1):
                if not self.cv.wait(timeout=0.05) and time.time() > deadline:
                    raise TimeoutError("mbarrier wait timed out (deadlock)")


class Model:
    NSTAGE = 3

    def __init__(self, iters, seed, ncomp=3, bug=None):
        self.iters, self.bug, self.nc = iters, bug, ncomp
        self.rng = random.Random(seed)
        self.deadline = time.time() + 20
        self.kv_full = MBar(1)
        self.qdo_full = [MBar(1) for _ in range(self.NSTAGE)]
        self.qdo_empty = [MBar(1) for _ in range(self.NSTAGE)]
        self.st_full = [MBar(1), MBar(1)]
        self.dp_full, self.dq_full, self.acc_full = MBar(1), MBar(1), MBar(1)
        self.pt_ready, self.dst_ready, self.dqt_empty = MBar(ncomp), MBar(ncomp), MBar(ncomp)
        # tagged resources: value = tile index held, readers outstanding
        self.QdO = [None] * self.NSTAGE
        self.ST, self.ST_readers = [None, None], [0, 0]
        self.DP, self.DP_readers = None, 0
        self.DQ, self.DQ_readers = None, 0
        self.PT, self.PT_unread = None, False
        self.DST, self.DST_unread = None, 0
        self.pipe = queue.Queue()
        self.errors = []
        self.lock = threading.Lock()

    def nap(self, scale=1.0):
        time.sleep(self.rng.random() * 0.001 * scale)

    def fail(self, msg):
        self.errors.append(msg)

    def pipe_thread(self):
        while True:
            op = self.pipe.get()
            if op is None:
                return
            op()

    def tma_thread(self):
        self.kv_full.arrive()
        for it in range(self.iters):
            st, ph = it % self.NSTAGE, (it // self.NSTAGE) & 1
            self.qdo_empty[st].wait(ph ^ 1, self.deadline)
            if self.QdO[st] is not None and self.QdO[st] != it - self.NSTAGE:
                self.fail(f"Q/dO stage {st} reloaded over tile {self.QdO[st]} for tile {it}")
            self.nap(2.0)
            self.QdO[st] = it
            self.qdo_full[st].arrive()

    # ---- MMA issuer (program order == tensor-pipe order) ----
    def mma_thread(self):
        def S(it):
            s, st = it & 1, it % self.NSTAGE

            def run():
                self.nap()
                if self.QdO[st] != it:
                    self.fail(f"S^T({it}) read Q of tile {self.QdO[st]}")
                if self.ST_readers[s]:
                    self.fail(f"S^T({it}) overwrote ST[{s}] (tile {self.ST[s]}) with readers outstanding")
                self.ST[s], self.ST_readers[s] = it, self.nc
            self.pipe.put(run)
            self.pipe.put(self.st_full[s].arrive)

        def dP(it):
            st = it % self.NSTAGE

            def run():
                self.nap()
                if self.QdO[st] != it:
                    self.fail(f"dP^T({it}) read dO of tile {self.QdO[st]}")
                if self.DP_readers:
                    self.fail(f"dP^T({it}) overwrote DP (tile {self.DP}) with readers outstanding")
                self.DP, self.DP_readers = it, self.nc
            self.pipe.put(run)
            self.pipe.put(self.dp_full.arrive)
        self.kv_full.wait(0, self.deadline)
        self.qdo_full[0].wait(0, self.deadline)
        S(0)
        dP(0)
        for it in range(self.iters):
            st, par = it % self.NSTAGE, it & 1
            self.pt_ready.wait(par, self.deadline)

            def dV(it=it, st=st):
                self.nap()
                if self.PT != it or self.QdO[st] != it:
                    self.fail(f"dV({it}) read P^T of tile {self.PT} / dO of tile {self.QdO[st]}")
                self.PT_unread = False
            self.pipe.put(dV)
            if it + 1 < self.iters:
                st1 = (it + 1) % self.NSTAGE
                self.qdo_full[st1].wait(((it + 1) // self.NSTAGE) & 1, self.deadline)
                S(it + 1)
            if self.bug != "dp_before_dst_ready":
                self.dst_ready.wait(par, self.deadline)
            if it + 1 < self.iters:
                dP(it + 1)
            if self.bug == "dp_before_dst_ready":
                self.dst_ready.wait(par, self.deadline)

            def dK(it=it, st=st):
                self.nap()
                if self.DST != it or self.QdO[st] != it:
                    self.fail(f"dK({it}) read dS^T of tile {self.DST} / Q of tile {self.QdO[st]}")
                self.DST_unread -= 1
            self.pipe.put(dK)
            if self.bug != "no_dqt_empty":
                self.dqt_empty.wait(par ^ 1, self.deadline)

            def dQ(it=it):
                self.nap()
                if self.DST != it:
                    self.fail(f"dQ({it}) read dS^T of tile {self.DST}")
                if self.DQ_readers:
                    self.fail(f"dQ^T({it}) overwrote DQ (tile {self.DQ}) before staging read it")
                self.DQ, self.DQ_readers = it, self.nc
                self.DST_unread -= 1
            self.pipe.put(dQ)
            self.pipe.put(self.qdo_empty[st].arrive)
            self.pipe.put(self.dq_full.arrive)
        self.pipe.put(self.acc_full.arrive)

    # ---- compute warps ----
    def staging(self, t, k):
        """dQ^T of tile k: TMEM -> staging -> bulk reduce (the caller has already waited dq_full for tile k)."""
        self.nap(4.0)                                   # adversarial: staging is slow relative to the tensor pipe
        if self.DQ != k:
            self.fail(f"staging({k}) read dQ^T of tile {self.DQ}")
        with self.lock:
            self.DQ_readers -= 1
        self.dqt_empty.arrive()

    def compute_thread(self, t):
        for it in range(self.iters):
            s, ph, par = it & 1, (it >> 1) & 1, it & 1
            self.st_full[s].wait(ph, self.deadline)
            if self.ST[s] != it:
                self.fail(f"P({it}) read S^T of tile {self.ST[s]}")
            self.nap()
            with self.lock:
                self.ST_readers[s] -= 1
            # P^T smem is free without a barrier of its own: S^T(it) was issued after dV(it-1) on the in-order
            # tensor pipe, so st_full(it) implies dV(it-1) has finished reading it (the model checks that claim)
            if t == 0:
                if self.PT_unread:
                    self.fail(f"P^T of tile {self.PT} overwritten by tile {it} before dV read it")
                self.PT = it
            self.nap()
            self._arrive_last(self.pt_ready, lambda: setattr(self, "PT_unread", True))
            self.dp_full.wait(par, self.deadline)
            self.nap(4.0)                               # adversarial: reading dP^T out of TMEM is slow
            if self.DP != it:
                self.fail(f"dS({it}) read dP^T of tile {self.DP}")
            with self.lock:
                self.DP_readers -= 1
            if it >= 1:
                parity = (it - 1) & 1
                if self.bug == "dq_full_parity":
                    parity ^= 1
                self.dq_full.wait(parity, self.deadline)                # dS^T smem free: dK/dQ(it-1) have read it
            if t == 0:
                if self.DST_unread:
                    self.fail(f"dS^T of tile {self.DST} overwritten by tile {it} with {self.DST_unread} GEMMs to go")
                self.DST = it
            self.nap()
            self._arrive_last(self.dst_ready, lambda: setattr(self, "DST_unread", 2))
            if it >= 1:
                self.staging(t, it - 1)
        self.dq_full.wait((self.iters - 1) & 1, self.deadline)
        self.staging(t, self.iters - 1)
        self.acc_full.wait(0, self.deadline)

    def _arrive_last(self, bar, on_last):
        with bar.cv:
            last = bar.pending == 1
        if last:
            on_last()
        bar.arrive()

    def _guard(self, fn):
        try:
            fn()
        except Exception as exc:  # noqa: BLE001
            self.fail(repr(exc))

    def run(self):
        fns = [self.pipe_thread, self.tma_thread, self.mma_thread] + [lambda t=t: self.compute_thread(t) for t in range(self.nc)]
        threads = [threading.Thread(target=self._guard, args=(fn,), daemon=True) for fn in fns]
        for th in threads:
            th.start()
        for th in threads[1:]:
            th.join(30)
        self.pipe.put(None)
        threads[0].join(5)
        if any(th.is_alive() for th in threads):
            self.fail("model deadlocked")
        return self.errors


@pytest.mark.parametrize("iters", [1, 2, 3, 4, 5, 8, 13])
def test_bwd_v3_protocol_is_safe_and_live(iters):
    for seed in range(4):
        assert Model(iters, seed).run() == [], (iters, seed)


@pytest.mark.parametrize("bug", ["dq_full_parity", "no_dqt_empty", "dp_before_dst_ready"])
def test_model_catches_protocol_mistakes(bug):
    caught = False
    for seed in range(10):
        m = Model(9, seed, bug=bug)
        m.deadline = time.time() + 3
        if m.run():
            caught = True
            break
    assert caught, bug
