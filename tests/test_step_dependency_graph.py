"""Is a circular wait possible between the kernels of one ZeRO-3 backward pass with the fused collectives?

`test_flag_protocol_model.py` runs the flag protocol with threads and random delays. This file asks the static question
for the engine's actual launch structure (models/native_llama.py + engine/zero.py + parallel/fused.py): every kernel
instance of every rank is a node, every reason one of them has to wait for another — stream order, CUDA events between
the two streams, the two pooled gradient buffers, the pooled parameter buffers, the cross-GPU flags of K1 — is an edge.
If that graph is acyclic, no timing of launches can deadlock the step as long as every launched kernel eventually
gets an SM (occupancy is a separate matter, see DESIGN.md "The 8-GPU stall"). Two negative tests show that the check
would catch a wrong design."""
import itertools

import pytest


def build_graph(ws: int, units: int, param_slots: int = 4, grad_slots: int = 2, gathers_wait_for_reduce: bool = True,
                ready_waits_for_own_reduce_end: bool = False):
    """Backward over units U..1 (+ root unit 0 at the end). Per rank r and unit u:
        bwd(r,u)     main stream: the block's backward kernels
        gather(r,u)  comm stream: pull all-gather of the block's weights (only blocks not resident from forward)
        ready(r,u)   comm stream: K1 phase 0 — announce "my bucket of u is complete", wait for everybody's announcement
        reduce(r,u)  comm stream: K1 reads every rank's bucket
        done(r,u)    comm stream: K1 phase 1 — announce "I have read", wait for everybody's; the bucket may be reused
    Returns {node: set(prerequisite nodes)}."""
    g = {}

    def need(node, *pre):
        g.setdefault(node, set()).update(p for p in pre if p is not None)

    order = list(range(units, 0, -1)) + [0]
    resident_after_forward = set(range(units, max(units - param_slots, 0), -1))
    for r in range(ws):
        main_prev, comm_prev = None, None
        for i, u in enumerate(order):
            # --- comm stream: prefetch of the NEXT block's weights is issued first (engine order) ---
            nxt = order[i + 1] if i + 1 < len(order) else None
            if nxt is not None and nxt >= 1 and nxt not in resident_after_forward:
                gat = ("gather", r, nxt)
                # its parameter buffer held unit nxt + param_slots, whose backward is over by now on the main stream
                need(gat, comm_prev if gathers_wait_for_reduce else None, main_prev)
                comm_prev = gat if gathers_wait_for_reduce else comm_prev
            # --- main stream: backward of unit u ---
            bwd = ("bwd", r, u)
            need(bwd, main_prev)
            if u >= 1 and u not in resident_after_forward:
                need(bwd, ("gather", r, u))
            if u >= 1 and u + grad_slots <= units:
                need(bwd, ("done", r, u + grad_slots))           # the gradient bucket it writes into is free again
            main_prev = bwd
            # --- comm stream: K1 of unit u ---
            ready, red, done = ("ready", r, u), ("reduce", r, u), ("done", r, u)
            need(ready, comm_prev, bwd)
            for q in range(ws):                                   # phase 0: every rank's announcement
                need(ready, ("launched", q, u))
            need(("launched", r, u), comm_prev, bwd)              # a rank announces when its ready kernel starts
            if ready_waits_for_own_reduce_end:
                need(("launched", r, u), done)                    # (a wrong design, for the negative test)
            need(red, ready)
            need(done, red)
            for q in range(ws):                                   # phase 1: every rank has finished reading
                need(done, ("reduce", q, u))
            comm_prev = done
        need(("step", r), main_prev, comm_prev)                   # main stream waits for the comm stream, then all-reduce
    for r, q in itertools.product(range(ws), repeat=2):
        need(("optimizer", r), ("step", q))                       # the control all-reduce joins all ranks
    return g


def find_cycle(g):
    WHITE, GREY, BLACK = 0, 1, 2
    colour = {}
    for root in g:
        if colour.get(root, WHITE) != WHITE:
            continue
        stack = [(root, iter(g.get(root, ())))]
        colour[root] = GREY
        path = [root]
        while stack:
            node, it = stack[-1]
            for nxt in it:
                c = colour.get(nxt, WHITE)
                if c == GREY:
                    return path[path.index(nxt):] + [nxt]
                if c == WHITE:
                    colour[nxt] = GREY
                    stack.append((nxt, iter(g.get(nxt, ()))))
                    path.append(nxt)
                    break
            else:
                colour[node] = BLACK
                stack.pop()
                path.pop()
    return None


@pytest.mark.parametrize("ws", [2, 4, 8])
@pytest.mark.parametrize("units,param_slots", [(6, 2), (6, 4), (32, 4)])
def test_backward_with_fused_collectives_has_no_circular_wait(ws, units, param_slots):
    g = build_graph(ws, units, param_slots)
    assert find_cycle(g) is None
    assert find_cycle(build_graph(ws, units, param_slots, gathers_wait_for_reduce=False)) is None


def test_check_has_teeth_announce_after_own_phase_one():
    """If a rank only announced its bucket once its own K1 had finished, everybody would wait for everybody."""
    cyc = find_cycle(build_graph(2, 4, ready_waits_for_own_reduce_end=True))
    assert cyc is not None and any(n[0] == "launched" for n in cyc)


def test_check_has_teeth_single_gradient_bucket_reused_too_early():
    """With one pooled gradient bucket, a block's backward would need the reduce of the block before it to be over —
    which it is (comm stream order), so that is still fine; but asking for the reduce of the SAME unit first is a cycle."""
    g = build_graph(2, 4)
    g[("bwd", 0, 2)].add(("done", 0, 2))
    assert find_cycle(g) is not None
