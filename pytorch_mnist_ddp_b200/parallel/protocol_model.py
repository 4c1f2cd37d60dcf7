"""Happens-before model of the cross-GPU protocol of one training step (race / deadlock check on the CPU).

The data-parallel kernels of csrc/allreduce.cu end WITHOUT exit barriers: that a peer never reads a gradient, a
reduced chunk or a parameter slice too early or too late rests on an ordering argument over three kernels per rank,
two branch streams and five kinds of flag words.  compute-sanitizer cannot see across GPUs (and is not available on
the pool this was developed on), so the argument is checked here mechanically instead of by hand.

The model: every rank executes, per step, the kernels of csrc/engine.cu:enqueue_train as a DAG of events

    W(cell, version)    a store that makes `cell` hold the value of `version`
    R(cell, version)    a load that must observe exactly `version`
    SIG(slot, token)    st.release.sys of a flag word        (comm_barrier / comm_signal_owner / "R complete")
    WAIT(slot, token)   ld.acquire.sys spin until slot >= token

with program order inside a CTA, stream/event order between kernels of one rank, and one synchronisation edge
SIG(slot, tok) -> WAIT(slot, tok) (tokens grow monotonically, 2*step+1, and every slot has a single writer, so the
first store that satisfies a wait is the one carrying exactly its token).  All waits are unconditional, so the set of
executions is exactly the set of linearisations of this graph, and the checks are exact for the model:

  * the graph is acyclic                                  -> no deadlock (a wait that needs a later signal)
  * the stores to one cell are totally ordered            -> no write/write race
  * every load is ordered after the store of the version it expects and before the next store to that cell
                                                          -> no stale read, no early overwrite
  * the signals of one slot are totally ordered           -> the token argument above holds

What is modelled (file:line = where the real thing is):
  fused_allreduce_kernel  csrc/allreduce.cu  (fc bucket; one-shot / two-shot; NVLS has the two-shot protocol)
  conv2_reduce_kernel     csrc/allreduce.cu  (owner = chunk % W gathers, all-reduces into every rank's R, raises
                                              "R complete"; one-shot: all-to-all barrier per chunk)
  conv_finalize_kernel    csrc/allreduce.cu  (conv2 CTAs poll the local flag; conv1 CTAs all-to-all barrier)
  step graph              csrc/engine.cu:enqueue_train (main stream, fc branch, conv2 branch, the joins)
  rank-0-only work between steps (evaluation / checkpoint: reads every parameter; mnist_ddp.py:183-189)
  b200_comm_run back to back  (StandaloneModel: the stand-alone entry point, where the exit barrier IS needed)

`mutate=` removes one ordering at a time; tests/test_protocol_model.py asserts that each removal is caught, which is
what makes a clean result on the unmutated model mean something.
"""
from __future__ import annotations

import argparse
from collections import defaultdict
from typing import Dict, List, Optional, Tuple

ONESHOT, TWOSHOT = 0, 1

MUTATIONS = (
    "no_finalize_barrier",   # conv1 chunks exchanged without the step's all-to-all barrier
    "no_r_complete_wait",    # finalize reads R without polling the owner's flag
    "no_r_complete_signal_order",  # owner raises "R complete" before storing R
    "fc_before_dgrad",       # fc-bucket kernel not ordered after fc1 dgrad (cudaStreamWaitEvent(ev_fcd) dropped)
    "no_join_fc",            # finalize not ordered after the fc branch (ev_fcu join dropped)
    "no_join_conv2",         # finalize not ordered after the conv2 branch (ev_c2w join dropped)
    "no_gather_wait",        # conv2 owner reads the peers' chunks without waiting for their signals
    "no_fc_entry_barrier",   # fc-bucket kernel reads the peers' gradients without the entry barrier
    "stale_token",           # barrier tokens do not advance from step to step
)


class Violation(Exception):
    pass


class Graph:
    """Events + edges; reachability by bitsets over a topological order."""

    def __init__(self) -> None:
        self.kind: List[str] = []
        self.obj: List[tuple] = []
        self.val: List[int] = []
        self.rank: List[int] = []
        self.label: List[str] = []
        self.succ: List[List[int]] = []

    def add(self, rank: int, kind: str, obj: tuple, val: int, label: str, after: Tuple[int, ...] = ()) -> int:
        i = len(self.kind)
        self.kind.append(kind); self.obj.append(obj); self.val.append(val); self.rank.append(rank)
        self.label.append(label); self.succ.append([])
        for a in after:
            self.edge(a, i)
        return i

    def edge(self, a: int, b: int) -> None:
        self.succ[a].append(b)

    def describe(self, i: int) -> str:
        return f"rank {self.rank[i]} {self.label[i]}: {self.kind[i]}({self.obj[i]}, {self.val[i]})"

    def closure(self) -> List[int]:
        """reach[i] = bitset of events that happen after i (raises on a cycle = deadlock)."""
        n = len(self.kind)
        indeg = [0] * n
        for a in range(n):
            for b in self.succ[a]:
                indeg[b] += 1
        order, stack = [], [i for i in range(n) if indeg[i] == 0]
        while stack:
            a = stack.pop()
            order.append(a)
            for b in self.succ[a]:
                indeg[b] -= 1
                if indeg[b] == 0:
                    stack.append(b)
        if len(order) != n:
            stuck = [i for i in range(n) if indeg[i] > 0][:4]
            raise Violation("deadlock: cyclic wait through " + "; ".join(self.describe(i) for i in stuck))
        reach = [0] * n
        for a in reversed(order):
            m = 0
            for b in self.succ[a]:
                m |= reach[b] | (1 << b)
            reach[a] = m
        return reach


class StepModel:
    """Builds the event graph of `steps` training steps on `world` ranks."""

    def __init__(self, world: int, steps: int, fc_mode: int = TWOSHOT, conv2_mode: int = TWOSHOT, chunks: int = 0,
                 fc_ctas: int = 2, eval_after: Optional[int] = 0, ordered_conv2: bool = True,
                 mutate: Optional[str] = None) -> None:
        if mutate is not None and mutate not in MUTATIONS:
            raise ValueError(f"unknown mutation {mutate}")
        self.W, self.S, self.fc_mode, self.c2_mode = world, steps, fc_mode, conv2_mode
        self.chunks = chunks or world + 1      # conv2 chunks: every owner has one, owner 0 has two
        self.fc_ctas = fc_ctas
        self.eval_after = eval_after           # rank 0 evaluates after this step (None: never)
        self.ordered = ordered_conv2
        self.mut = mutate
        self.g = Graph()
        self._build()

    # -- helpers ------------------------------------------------------------------------------------------
    def tok(self, t: int) -> int:
        return 1 if self.mut == "stale_token" else 2 * t + 1

    def seq(self, rank: int, label: str, ops: List[Tuple[str, tuple, int]], after: Tuple[int, ...]) -> Tuple[int, int]:
        """A chain of events in program order; returns (first, last)."""
        first = last = -1
        for kind, obj, val in ops:
            last = self.g.add(rank, kind, obj, val, label, after if first < 0 else (last,))
            if first < 0:
                first = last
        if first < 0:  # empty: a no-op node keeps the DAG simple
            first = last = self.g.add(rank, "NOP", (), 0, label, after)
        return first, last

    def kernel(self, rank: int, label: str, ctas: List[List[Tuple[str, tuple, int]]], after: Tuple[int, ...]) -> int:
        """A kernel: launch node -> independent CTAs -> completion node (what the next kernel in the stream waits for)."""
        start = self.g.add(rank, "NOP", (), 0, label + " launch", after)
        ends = [self.seq(rank, f"{label} cta{k}", ops, (start,))[1] for k, ops in enumerate(ctas)]
        return self.g.add(rank, "NOP", (), 0, label + " done", tuple(ends))

    # -- the three communication kernels ---------------------------------------------------------------------
    def fc_kernel(self, r: int, t: int) -> List[List[Tuple[str, tuple, int]]]:
        W, tk = self.W, self.tok(t)
        ctas = []
        for b in range(self.fc_ctas):
            ops: List[Tuple[str, tuple, int]] = [("R", ("step", r), t)]
            own = [r] if self.fc_mode == TWOSHOT else list(range(W))
            for o in own:  # state / parameter prefetch of the CTA's own elements, before the barrier
                ops.append(("R", ("Pfc", r, o, b), t))
            if self.mut != "no_fc_entry_barrier":
                ops += [("SIG", ("fc", b, r, q), tk) for q in range(W)]    # slot [b][r] in rank q's flag block
                ops += [("WAIT", ("fc", b, q, r), tk) for q in range(W)]   # my slots [b][q]
            for o in own:
                ops += [("R", ("Gfc", q, o, b), t + 1) for q in range(W)]
                dst = range(W) if self.fc_mode == TWOSHOT else [r]
                ops += [("W", ("Pfc", q, o, b), t + 1) for q in dst]
            ctas.append(ops)
        return ctas

    def conv2_kernel(self, r: int, t: int) -> List[List[Tuple[str, tuple, int]]]:
        W, tk = self.W, self.tok(t)
        ctas = []
        for c in range(self.chunks):
            ops: List[Tuple[str, tuple, int]] = [("R", ("step", r), t), ("R", ("c2part", r), t + 1),
                                                  ("W", ("Gc2", r, c), t + 1)]
            if self.c2_mode == ONESHOT:
                ops += [("SIG", ("c2", c, r, q), tk) for q in range(W)]
                ops += [("WAIT", ("c2", c, q, r), tk) for q in range(W)]
                ops += [("R", ("Gc2", q, c), t + 1) for q in range(W)]
                ops += [("W", ("R", r, c), t + 1), ("SIG", ("rdone", c, r, r), tk)]
            else:
                o = c % W
                if r != o:
                    ops.append(("SIG", ("c2", c, r, o), tk))
                else:
                    if self.mut != "no_gather_wait":
                        ops += [("WAIT", ("c2", c, q, o), tk) for q in range(W) if q != o]
                    ops += [("R", ("Gc2", q, c), t + 1) for q in range(W)]
                    stores = [("W", ("R", q, c), t + 1) for q in range(W)]
                    sigs = [("SIG", ("rdone", c, o, q), tk) for q in range(W)]
                    ops += (sigs + stores) if self.mut == "no_r_complete_signal_order" else (stores + sigs)
            ctas.append(ops)
        return ctas

    def finalize_kernel(self, r: int, t: int) -> List[List[Tuple[str, tuple, int]]]:
        W, tk = self.W, self.tok(t)
        ctas = []
        for c in range(self.chunks):  # conv2 chunks: local poll, read R, Adadelta on the chunk
            o = c % W if self.c2_mode != ONESHOT else r
            ops: List[Tuple[str, tuple, int]] = [("R", ("step", r), t)]
            if self.mut != "no_r_complete_wait":
                ops.append(("WAIT", ("rdone", c, o, r), tk))
            ops += [("R", ("R", r, c), t + 1), ("R", ("Pc2", r, c), t), ("W", ("Pc2", r, c), t + 1)]
            ctas.append(ops)
        ops = [("R", ("step", r), t), ("R", ("c1part", r), t + 1), ("W", ("Gc1", r), t + 1)]
        if self.mut != "no_finalize_barrier":
            ops += [("SIG", ("fin", 0, r, q), tk) for q in range(W)]
            ops += [("WAIT", ("fin", 0, q, r), tk) for q in range(W)]
        ops += [("R", ("Gc1", q), t + 1) for q in range(W)]
        ops += [("R", ("Pc1", r), t), ("W", ("Pc1", r), t + 1)]
        ctas.append(ops)
        return ctas

    # -- one rank's step graph (csrc/engine.cu:enqueue_train) -------------------------------------------------
    def _params_read(self, r: int, ver: int) -> List[Tuple[str, tuple, int]]:
        ops = [("R", ("Pfc", r, o, b), ver) for o in range(self.W) for b in range(self.fc_ctas)]
        ops += [("R", ("Pc2", r, c), ver) for c in range(self.chunks)]
        ops.append(("R", ("Pc1", r), ver))
        return ops

    def _build(self) -> None:
        g, W = self.g, self.W
        init = []
        for r in range(W):  # version 0 of every cell, before anything else
            ops = [("W", ("Pfc", r, o, b), 0) for o in range(W) for b in range(self.fc_ctas)]
            ops += [("W", ("Gfc", r, o, b), 0) for o in range(W) for b in range(self.fc_ctas)]
            ops += [("W", ("Pc2", r, c), 0) for c in range(self.chunks)] + [("W", ("Gc2", r, c), 0) for c in range(self.chunks)]
            ops += [("W", ("R", r, c), 0) for c in range(self.chunks)]
            ops += [("W", ("Pc1", r), 0), ("W", ("Gc1", r), 0), ("W", ("c2part", r), 0), ("W", ("c1part", r), 0),
                    ("W", ("step", r), 0)]
            init.append(self.seq(r, "init", ops, ())[1])
        start = g.add(0, "NOP", (), 0, "rendezvous (attach(): host barrier after the parameter pull)", tuple(init))
        main = [start] * W
        br0: List[Tuple[int, ...]] = [()] * W   # last kernel on the fc branch / the conv2 branch: eager launches keep
        br1: List[Tuple[int, ...]] = [()] * W   # stream order there even where a join is mutated away
        for t in range(self.S):
            for r in range(W):
                m = main[r]
                # forward (conv12, fc1, head): reads every parameter
                m = self.kernel(r, f"step {t} forward", [self._params_read(r, t)], (m,))
                ev_head = m
                # main: fc1 dgrad reads fc1.weight
                dgrad = self.kernel(r, f"step {t} fc1 dgrad",
                                    [[("R", ("Pfc", r, o, b), t) for o in range(W) for b in range(self.fc_ctas)]], (m,))
                # fc branch: head wgrad + fc1 wgrad write the fc gradient bucket, then the fused kernel
                wg = self.kernel(r, f"step {t} fc wgrad",
                                 [[("W", ("Gfc", r, o, b), t + 1) for o in range(W) for b in range(self.fc_ctas)]], (ev_head,) + br0[r])
                fc_after = (wg,) if self.mut == "fc_before_dgrad" else (wg, dgrad)
                fc = self.kernel(r, f"step {t} fc bucket", self.fc_kernel(r, t), fc_after)
                # main: conv2 wgrad -> (fork) conv2 reduce on branch 1; conv2 dgrad; conv1 wgrad
                c2w = self.kernel(r, f"step {t} conv2 wgrad", [[("W", ("c2part", r), t + 1)]], (dgrad,))
                c2r = self.kernel(r, f"step {t} conv2 reduce", self.conv2_kernel(r, t), (c2w,) + br1[r])
                br0[r], br1[r] = (fc,), (c2r,)
                c2d_after = (c2w,) if self.ordered else (dgrad,)
                c2d = self.kernel(r, f"step {t} conv2 dgrad", [[("R", ("Pc2", r, c), t) for c in range(self.chunks)]], c2d_after)
                c1w = self.kernel(r, f"step {t} conv1 wgrad", [[("W", ("c1part", r), t + 1)]], (c2d,))
                join = [c1w]
                if self.mut != "no_join_fc":
                    join.append(fc)
                if self.mut != "no_join_conv2":
                    join.append(c2r)
                fin = self.kernel(r, f"step {t} finalize", self.finalize_kernel(r, t), tuple(join))
                m = g.add(r, "W", ("step", r), t + 1, f"step {t} finalize: last CTA advances the step counter", (fin,))
                if r == 0 and self.eval_after == t:  # rank-0-only work between steps
                    m = self.kernel(0, f"evaluation after step {t}", [self._params_read(0, t + 1)], (m,))
                main[r] = m

    # -- the checks --------------------------------------------------------------------------------------------
    def check(self) -> Dict[str, int]:
        g = self.g
        sigs: Dict[tuple, List[int]] = defaultdict(list)
        for i, k in enumerate(g.kind):
            if k == "SIG":
                sigs[g.obj[i]].append(i)
        # synchronisation edges: the wait completes on the first store whose token is >= the awaited one.  Stores to a
        # slot come from one CTA of one rank, kernel after kernel; order them by token for the search and verify below
        # that this IS their execution order.
        n_wait = 0
        for i, k in enumerate(g.kind):
            if k != "WAIT":
                continue
            n_wait += 1
            cands = [j for j in sigs.get(g.obj[i], []) if g.val[j] >= g.val[i]]
            if not cands:
                raise Violation("deadlock: nobody ever signals " + g.describe(i))
            # with monotone tokens the earliest satisfying store is the one with the smallest token; with equal tokens
            # (a protocol bug) ANY of them may satisfy it, i.e. only the first issued is guaranteed
            first = min(cands, key=lambda j: (g.val[j], j))
            g.edge(first, i)
        reach = g.closure()
        hb = lambda a, b: bool((reach[a] >> b) & 1)
        for slot, lst in sigs.items():
            lst.sort(key=lambda j: (g.val[j], j))
            for a, b in zip(lst, lst[1:]):
                if g.val[a] == g.val[b]:
                    raise Violation(f"token reuse on slot {slot}: {g.describe(a)} and {g.describe(b)} -- a wait for the "
                                    "second may be satisfied by the first")
                if not hb(a, b):
                    raise Violation(f"signals on slot {slot} are unordered: {g.describe(a)} / {g.describe(b)}")
        writes: Dict[tuple, List[int]] = defaultdict(list)
        for i, k in enumerate(g.kind):
            if k == "W":
                writes[g.obj[i]].append(i)
        for cell, lst in writes.items():
            lst.sort(key=lambda j: g.val[j])
            for a, b in zip(lst, lst[1:]):
                if not hb(a, b):
                    raise Violation(f"write/write race on {cell}: {g.describe(a)} / {g.describe(b)}")
        n_read = 0
        for i, k in enumerate(g.kind):
            if k != "R":
                continue
            n_read += 1
            lst = writes.get(g.obj[i], [])
            src = [j for j in lst if g.val[j] == g.val[i]]
            if not src:
                raise Violation("no producer for " + g.describe(i))
            if not hb(src[0], i):
                raise Violation(f"stale read: {g.describe(i)} is not ordered after {g.describe(src[0])}")
            nxt = [j for j in lst if g.val[j] > g.val[i]]
            if nxt and not hb(i, nxt[0]):
                raise Violation(f"early overwrite: {g.describe(nxt[0])} is not ordered after {g.describe(i)}")
        return {"events": len(g.kind), "waits": n_wait, "reads": n_read, "cells": len(writes), "slots": len(sigs)}


class StandaloneModel(StepModel):
    """`b200_comm_run` called back to back on one bucket (tests, benchmarks/allreduce_bench.py): the caller refills the
    gradient buffer between calls, nothing else orders the ranks, so here the kernel's EXIT barrier is what keeps a
    fast rank's refill away from a slow rank's loads.  Tokens grow by 2 per call; the exit barrier uses token + 1 on
    the phase-1 slots."""

    def __init__(self, world: int, calls: int, mode: int = TWOSHOT, exit_barrier: bool = True, ctas: int = 2) -> None:
        self.W, self.S, self.fc_mode, self.fc_ctas, self.mut = world, calls, mode, ctas, None
        self.exit_barrier = exit_barrier
        self.g = Graph()
        g, W = self.g, world
        init = []
        for r in range(W):
            ops = [("W", ("P", r, o, b), 0) for o in range(W) for b in range(ctas)]
            ops += [("W", ("G", r, o, b), 0) for o in range(W) for b in range(ctas)]
            init.append(self.seq(r, "init", ops, ())[1])
        start = g.add(0, "NOP", (), 0, "rendezvous", tuple(init))
        last = [start] * W
        for t in range(calls):
            for r in range(W):
                fill = self.kernel(r, f"call {t} fill gradients",
                                   [[("W", ("G", r, o, b), t + 1) for o in range(W) for b in range(ctas)]], (last[r],))
                last[r] = self.kernel(r, f"call {t} all-reduce", self._run(r, t), (fill,))

    def _run(self, r: int, t: int):
        W, tk = self.W, 2 * t + 1
        ctas = []
        for b in range(self.fc_ctas):
            ops = [("SIG", ("p0", b, r, q), tk) for q in range(W)] + [("WAIT", ("p0", b, q, r), tk) for q in range(W)]
            own = [r] if self.fc_mode == TWOSHOT else list(range(W))
            for o in own:
                ops += [("R", ("G", q, o, b), t + 1) for q in range(W)]
                ops += [("W", ("P", q, o, b), t + 1) for q in (range(W) if self.fc_mode == TWOSHOT else [r])]
            if self.exit_barrier:
                ops += [("SIG", ("p1", b, r, q), tk + 1) for q in range(W)]
                ops += [("WAIT", ("p1", b, q, r), tk + 1) for q in range(W)]
            ctas.append(ops)
        return ctas


def verify_standalone(world: int, calls: int = 3, mode: int = TWOSHOT, exit_barrier: bool = True) -> Dict[str, int]:
    return StandaloneModel(world, calls, mode, exit_barrier).check()


def verify(world: int, steps: int = 3, fc_mode: int = TWOSHOT, conv2_mode: int = TWOSHOT, mutate: Optional[str] = None,
           **kw) -> Dict[str, int]:
    """Raises Violation with the two offending events, or returns the size of what was checked."""
    return StepModel(world, steps, fc_mode, conv2_mode, mutate=mutate, **kw).check()


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--mutate", choices=MUTATIONS, default=None)
    a = ap.parse_args()
    for fc in (ONESHOT, TWOSHOT):
        for c2 in (ONESHOT, TWOSHOT):
            for ordered in (True, False):
                tag = f"world {a.world}, fc {'one' if fc == 0 else 'two'}-shot, conv2 {'one' if c2 == 0 else 'two'}-shot, " \
                      f"conv2 wgrad/dgrad {'ordered' if ordered else 'unordered'}"
                try:
                    st = verify(a.world, a.steps, fc, c2, a.mutate, ordered_conv2=ordered)
                    print(f"ok    {tag}: {st}")
                except Violation as e:
                    print(f"FAIL  {tag}: {e}")


if __name__ == "__main__":
    main()
