"""Litmus model of the frame kernels' hand-off protocol (csrc/ocean_frame.cu, DESIGN.md "Frame scheduling").

compute-sanitizer is closed on the GPU pool this library is developed on (profiles/r2_sanitizer.txt), so the protocol the
two frame kernels synchronise with -- hand-rolled release/acquire counters between co-resident grids, a proxy fence in
front of the TMA reads, bounded spins -- is checked here instead: an abstract machine small enough to enumerate EVERY
interleaving, with the same operations in the same order as the kernels, and a weak enough memory to expose a missing
fence.  Like fft_plan.py this is a restatement for the tests (tests/test_handoff_model.py); nothing on the product path
imports it.

The machine
  memory    data cells (workspace rows, output tiles) and counters.  A data store goes into the issuing agent's store
            buffer and reaches memory when that agent fences (`fence`: __threadfence / cp.async.bulk.wait_group + fences) or
            whenever the scheduler decides to drain it (stores may become visible at any time, in order).  Counter updates
            are atomics: they act on memory at once and are NOT ordered after the agent's buffered stores unless it fenced.
  proxies   a TMA read (async proxy) of cells written through the generic proxy is only legal after the reading agent has
            executed `proxy_fence` since its last acquire (fence.proxy.async on the consumer side of the chain).
  agents    per frame f = 1..F:  row CTAs (layer, row) of ocean_rows_kernel and persistent column CTAs of
            ocean_cols_tma_kernel, each working through its tiles in order.
  launches  kernels are launched in stream order R(1) C(1) R(2) C(2) ...; every kernel triggers its dependents at once
            (griddepcontrol.launch_dependents at the top, no griddepcontrol.wait), so a CTA of kernel k+1 may start as soon
            as EVERY CTA of kernel k has started -- and a CTA starts only when the SM has room (`capacity`, `row_cost`,
            `col_cost`): this is what makes spinning consumers safe, and what the deadlock check explores.

The programs (one step = one transition; `ws[l][r]`, `out[l][c]` hold the frame number that last wrote them)
  row CTA (f, l, r):    start -> wait colsDone[l] >= K*(f-1) -> store ws[l][r] := f -> fence -> rowsDone[l] += 1 -> exit
  column CTA (f, j):    start -> for each of its tiles (l, c):
                            wait rowsDone[l] >= R*f -> proxy_fence -> TMA-read ws[l][*]   (must all read f)
                            -> TMA-read out[l][c]                                          (must read f-1: the foam chain)
                            -> store out[l][c] := f -> fence -> colsDone[l] += 1
                        -> exit
Checked in every reachable state: the two reads above, "a row store never lands on a cell a column tile of the previous
frame has yet to read" (follows from the first read being checked for every tile), and absence of deadlock (a state with
unfinished agents and no enabled transition).  `mutate` removes one ingredient at a time; the tests require the search to
find a violation for each.
"""
from collections import namedtuple

Config = namedtuple("Config", "frames layers rows tiles col_ctas capacity row_cost col_cost")
DEFAULT = Config(frames=2, layers=1, rows=2, tiles=2, col_ctas=1, capacity=3, row_cost=1, col_cost=2)

MUTATIONS = (
    "no_row_fence",         # rows publish without __threadfence: a column tile may read a stale workspace row
    "no_war_wait",          # rows do not wait for colsDone: frame f+1 overwrites rows frame f's columns still need
    "no_proxy_fence",       # the producer lane skips fence.proxy.async after acquiring the row counter
    "no_col_fence",         # columns publish before their stores completed: the next frame reads stale turbulence
    "no_row_wait",          # the column pass does not wait for its layer's rows at all
    "no_launch_rule",       # CTAs of kernel k+1 may start before every CTA of kernel k has started (no PDL guarantee)
)


class Violation(Exception):
    pass


def _programs(cfg, mutate):
    """agent id -> (kernel ordinal, cost, list of ops).  An op is a tuple; see step()."""
    agents = []
    for f in range(1, cfg.frames + 1):
        kr, kc = 2 * (f - 1), 2 * (f - 1) + 1
        for l in range(cfg.layers):
            for r in range(cfg.rows):
                ops = [("start",)]
                if mutate != "no_war_wait":
                    ops.append(("wait", ("colsDone", l), cfg.tiles * (f - 1)))
                ops.append(("store", ("ws", l, r), f))
                if mutate != "no_row_fence":
                    ops.append(("fence",))
                ops += [("add", ("rowsDone", l)), ("exit",)]
                agents.append((kr, cfg.row_cost, ops))
        tiles = [(l, c) for l in range(cfg.layers) for c in range(cfg.tiles)]
        for j in range(cfg.col_ctas):
            ops = [("start",)]
            for (l, c) in tiles[j::cfg.col_ctas]:
                if mutate != "no_row_wait":
                    ops.append(("wait", ("rowsDone", l), cfg.rows * f))
                if mutate != "no_proxy_fence":
                    ops.append(("proxy_fence",))
                for r in range(cfg.rows):
                    ops.append(("tma_read", ("ws", l, r), f))
                ops.append(("tma_read_async_written", ("out", l, c), f - 1))
                ops.append(("store", ("out", l, c), f))
                if mutate != "no_col_fence":
                    ops.append(("fence",))
                ops.append(("add", ("colsDone", l)))
            ops.append(("exit",))
            agents.append((kc, cfg.col_cost, ops))
    return agents


def explore(cfg=DEFAULT, mutate=None, max_states=2_000_000):
    """Exhaustive search over every interleaving.  Returns the number of distinct states visited; raises Violation with
    a description (and the trace that leads to it) on a bad read or a deadlock."""
    if mutate is not None and mutate not in MUTATIONS:
        raise ValueError(mutate)
    agents = _programs(cfg, mutate)
    n = len(agents)
    kernels = max(a[0] for a in agents) + 1
    by_kernel = [[i for i, a in enumerate(agents) if a[0] == k] for k in range(kernels)]
    cells = [("ws", l, r) for l in range(cfg.layers) for r in range(cfg.rows)]
    cells += [("out", l, c) for l in range(cfg.layers) for c in range(cfg.tiles)]
    cells += [("rowsDone", l) for l in range(cfg.layers)] + [("colsDone", l) for l in range(cfg.layers)]
    index = {c: i for i, c in enumerate(cells)}

    # state = (pcs, buffers, memory, proxy_ok)
    init = (tuple([0] * n), tuple([()] * n), tuple([0] * len(cells)), tuple([False] * n))
    seen = {init}
    stack = [(init, ())]

    def started(pcs, i):
        return pcs[i] >= 1

    def resident_cost(pcs):
        return sum(agents[i][1] for i in range(n) if 1 <= pcs[i] < len(agents[i][2]))

    def may_start(pcs, i):
        k = agents[i][0]
        if k > 0:
            prev = by_kernel[k - 1]
            if mutate == "no_launch_rule":
                if not any(started(pcs, j) for j in prev):
                    return False
            elif not all(started(pcs, j) for j in prev):
                return False
        return resident_cost(pcs) + agents[i][1] <= cfg.capacity

    while stack:
        (pcs, bufs, mem, pok), trace = stack.pop()
        succ = []
        for i in range(n):
            ops = agents[i][2]
            if pcs[i] >= len(ops):
                continue
            op = ops[pcs[i]]
            kind = op[0]
            npcs = pcs[:i] + (pcs[i] + 1,) + pcs[i + 1:]
            label = (i, op)
            if kind == "start":
                if may_start(pcs, i):
                    succ.append(((npcs, bufs, mem, pok), label))
            elif kind == "wait":
                if mem[index[op[1]]] >= op[2]:  # acquire: what follows sees memory; the async proxy must be fenced again
                    succ.append(((npcs, bufs, mem, pok[:i] + (False,) + pok[i + 1:]), label))
            elif kind == "store":
                nb = bufs[:i] + (bufs[i] + ((index[op[1]], op[2]),),) + bufs[i + 1:]
                succ.append(((npcs, nb, mem, pok), label))
            elif kind == "fence":
                m = list(mem)
                for a, v in bufs[i]:
                    m[a] = v
                succ.append(((npcs, bufs[:i] + ((),) + bufs[i + 1:], tuple(m), pok), label))
            elif kind == "add":
                a = index[op[1]]
                succ.append(((npcs, bufs, mem[:a] + (mem[a] + 1,) + mem[a + 1:], pok), label))
            elif kind == "proxy_fence":
                succ.append(((npcs, bufs, mem, pok[:i] + (True,) + pok[i + 1:]), label))
            elif kind in ("tma_read", "tma_read_async_written"):
                if kind == "tma_read" and not pok[i]:
                    raise Violation("agent %d: TMA read of %r without a proxy fence after the acquire; trace %r" % (i, op[1], trace + (label,)))
                got = mem[index[op[1]]]
                if got != op[2]:
                    raise Violation("agent %d read %r = %d, expected %d; trace %r" % (i, op[1], got, op[2], trace + (label,)))
                succ.append(((npcs, bufs, mem, pok), label))
            elif kind == "exit":
                if bufs[i]:  # stores of an exited CTA still drain; keep them in the buffer (drain transition below)
                    pass
                succ.append(((npcs, bufs, mem, pok), label))
            else:
                raise AssertionError(kind)
        for i in range(n):  # a buffered store may reach memory at any time (oldest first)
            if bufs[i]:
                (a, v), rest = bufs[i][0], bufs[i][1:]
                succ.append(((pcs, bufs[:i] + (rest,) + bufs[i + 1:], mem[:a] + (v,) + mem[a + 1:], pok), (i, ("drain", cells[a]))))
        if not succ and any(pcs[i] < len(agents[i][2]) for i in range(n)):
            stuck = [(i, agents[i][2][pcs[i]]) for i in range(n) if pcs[i] < len(agents[i][2])]
            raise Violation("deadlock: no agent can move; waiting: %r; trace %r" % (stuck, trace))
        for st, label in succ:
            if st not in seen:
                seen.add(st)
                if len(seen) > max_states:
                    raise RuntimeError("state space larger than max_states")
                stack.append((st, trace + (label,)))
    return len(seen)
