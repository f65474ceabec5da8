"""CPU model of the mbarrier protocol of the v3 flash-attention forward kernel (csrc/attention_v3.cu): the TMA
producer, the MMA issuer, the in-order tensor pipe and the softmax warps are threads; mbarriers have the hardware
semantics (arrival count, phase bit, ``wait(parity)`` passes iff the current phase bit differs from ``parity`` — so
a waiter that is two phases behind would alias). Buffers carry tags, and every read asserts that it sees the KV
block it expects and every write that the previous contents were consumed.

What v3 changes against v2, and what this model is here to check: P is double-buffered, ``pv_done`` is split per
buffer, the exp/pack phase only waits for the consumer of P_{j-2}, and only the lazy-rescale path waits for the
latest PV. The phase arithmetic of those waits is easy to get wrong and a mistake is a hang or a silent race on
the GPU. Negative tests corrupt one parity expression each and must be caught.
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
    def __init__(self, nblk, seed, n_softmax=4, bug=None):
        self.nblk, self.bug, self.nsm = nblk, bug, n_softmax
        self.rng = random.Random(seed)
        self.deadline = time.time() + 20
        mk = lambda c: [MBar(c), MBar(c)]  # noqa: E731
        self.q_full = MBar(1)
        self.k_full, self.k_empty, self.v_full, self.v_empty = mk(1), mk(1), mk(1), mk(1)
        self.s_full, self.s_empty = mk(1), mk(n_softmax)
        self.p_full, self.pv_done = mk(n_softmax), mk(1)
        self.K, self.V, self.S, self.P = [None, None], [None, None], [None, None], [None, None]
        self.S_unread, self.P_unread = [0, 0], [False, False]
        self.pv_completed = 0           # number of PV GEMMs the tensor pipe has finished
        self.pv_inflight = False
        self.pipe = queue.Queue()
        self.errors = []
        self.need = [[self.rng.random() < 0.35 or j == 0 for j in range(nblk)] for _ in range(n_softmax)]

    def nap(self, scale=1.0):
        time.sleep(self.rng.random() * 0.0015 * scale)

    def fail(self, msg):
        self.errors.append(msg)

    # ---- tensor pipe: executes queued ops strictly in order ----
    def pipe_thread(self):
        while True:
            op = self.pipe.get()
            if op is None:
                return
            op()

    def tma_thread(self):
        self.nap()
        self.q_full.arrive()
        for j in range(self.nblk):
            st, ph = j & 1, (j >> 1) & 1
            self.k_empty[st].wait(ph ^ 1, self.deadline)
            if self.K[st] is not None and self.K[st] != j - 2:
                self.fail(f"K[{st}] overwritten while holding block {self.K[st]} (loading {j})")
            self.nap()
            self.K[st] = j
            self.k_full[st].arrive()
            self.v_empty[st].wait(ph ^ 1, self.deadline)
            self.nap()
            self.V[st] = j
            self.v_full[st].arrive()

    def mma_thread(self):
        def issue_qk(j):
            st, ph = j & 1, (j >> 1) & 1
            self.k_full[st].wait(ph, self.deadline)
            self.s_empty[st].wait(ph ^ 1, self.deadline)

            def run():
                self.nap()
                if self.K[st] != j:
                    self.fail(f"QK({j}) read K block {self.K[st]}")
                if self.S_unread[st]:
                    self.fail(f"QK({j}) overwrote S[{st}] with {self.S_unread[st]} softmax readers outstanding")
                self.S[st], self.S_unread[st] = j, self.nsm
            self.pipe.put(run)
            self.pipe.put(self.k_empty[st].arrive)      # tcgen05.commit -> k_empty
            self.pipe.put(self.s_full[st].arrive)       # tcgen05.commit -> s_full
        self.q_full.wait(0, self.deadline)
        issue_qk(0)
        for j in range(self.nblk):
            if j + 1 < self.nblk:
                issue_qk(j + 1)
            st, ph = j & 1, (j >> 1) & 1
            self.v_full[st].wait(ph, self.deadline)
            self.p_full[st].wait(ph, self.deadline)

            def run(j=j, st=st):
                self.pv_inflight = True
                self.nap(2.0)
                if self.P[st] != j or self.V[st] != j:
                    self.fail(f"PV({j}) read P block {self.P[st]} / V block {self.V[st]}")
                self.P_unread[st] = False
                self.pv_completed += 1
                self.pv_inflight = False
            self.pipe.put(run)
            self.pipe.put(self.v_empty[st].arrive)
            self.pipe.put(self.pv_done[st].arrive)

    def softmax_thread(self, t):
        for j in range(self.nblk):
            st, ph = j & 1, (j >> 1) & 1
            self.s_full[st].wait(ph, self.deadline)
            if self.S[st] != j:
                self.fail(f"softmax({j}) read S block {self.S[st]}")
            self.nap()
            self.S_unread[st] -= 1
            self.s_empty[st].arrive()
            self.nap()
            if j >= 2:
                parity = ph if self.bug == "pbuf_parity" else ph ^ 1
                self.pv_done[st].wait(parity, self.deadline)
            if self.need[t][j] and j > 0:
                b = (j - 1) & 1
                parity = ((j - 1) >> 1) & 1
                if self.bug == "rescale_parity":
                    parity ^= 1
                if self.bug != "no_rescale_wait":
                    self.pv_done[b].wait(parity, self.deadline)
                if self.pv_inflight or self.pv_completed != j:
                    self.fail(f"rescale at block {j} with PV in flight / only {self.pv_completed} PVs complete")
                self.nap()
            # write P_j into buffer st: its previous contents (block j-2) must have been consumed
            if t == 0:
                if self.P_unread[st]:
                    self.fail(f"P[{st}] (block {self.P[st]}) overwritten by block {j} before its PV read it")
                self.P[st] = j
            self.nap()
            self.p_full[st].arrive()
            if t == 0:
                pass
        # epilogue waits for the last PV
        self.pv_done[(self.nblk - 1) & 1].wait(((self.nblk - 1) >> 1) & 1, self.deadline)
        if self.pv_completed != self.nblk:
            self.fail(f"epilogue read O after {self.pv_completed} of {self.nblk} PVs")

    def run(self):
        threads = [threading.Thread(target=self._guard, args=(fn,), daemon=True) for fn in
                   [self.pipe_thread, self.tma_thread, self.mma_thread] +
                   [lambda t=t: self.softmax_thread(t) for t in range(self.nsm)]]
        # mark P as unread when all softmax threads have arrived (phase completion); modelled by a watcher on p_full
        orig = [b.arrive for b in self.p_full]
        for st in (0, 1):
            def arrive(st=st):
                with self.p_full[st].cv:
                    last = self.p_full[st].pending == 1
                if last:
                    self.P_unread[st] = True
                orig[st]()
            self.p_full[st].arrive = arrive
        for th in threads:
            th.start()
        for th in threads[1:]:
            th.join(30)
        self.pipe.put(None)
        threads[0].join(5)
        if any(th.is_alive() for th in threads):
            self.fail("model deadlocked")
        return self.errors

    def _guard(self, fn):
        try:
            fn()
        except Exception as exc:  # noqa: BLE001
            self.fail(repr(exc))


@pytest.mark.parametrize("nblk", [1, 2, 3, 4, 7, 12])
def test_v3_protocol_is_safe_and_live(nblk):
    for seed in range(4):
        assert Model(nblk, seed).run() == [], (nblk, seed)


@pytest.mark.parametrize("bug", ["pbuf_parity", "rescale_parity", "no_rescale_wait"])
def test_model_catches_wrong_phase_arithmetic(bug):
    caught = False
    for seed in range(8):
        m = Model(9, seed, bug=bug)
        m.deadline = time.time() + 3
        if m.run():
            caught = True
            break
    assert caught, bug
