"""Exhaustive interleaving check of the mbarrier pipelines used by the tcgen05 kernels.

compute-sanitizer's racecheck is not available on the GPU pool, and it does not understand mbarrier
phase parity anyway.  The kernels here synchronise their warp roles with mbarriers that are waited on
by PARITY (``mbarrier.try_wait.parity``): one bit, so "the phase I want is complete" and "the phase two
before it is complete and nothing since" look the same.  A protocol is only sound if no waiter can
ever be asked that ambiguous question.  This module states each pipeline as a handful of small
sequential programs (producer, consumers) over a faithful model of the barrier, and explores EVERY
interleaving - including asynchronous completions (TMA loads, tcgen05.commit arrivals) landing in any
order - looking for

* a wait that succeeds before the data it guards has landed (``early``),
* a buffer overwritten while a reader that was granted it has not released it (``overwrite``),
* an arrival on a barrier whose phase is already complete (``over-arrival``),
* a state in which nobody can move before the programs are done (``deadlock``).

The collectives between GPUs are modelled the same way: the flag-in-payload ("LL") all-reduce
(stores that land whenever, polls for a tag; ``lost`` = the awaited tag was overwritten before it was
seen) and the barrier-based fused all-reduce + optimizer over successive steps (monotonic epochs in
the signal pads; ``stale`` = gradients or parameters read before or after the step they belong to).

It found nothing new - it was written AFTER the shortcut-ring race of the folded conv+BN GEMM
 had been found the hard way - but it reproduces that race in 40 states from
the protocol alone (``convbn_ring(two_full_barriers=False)``), passes the fixed protocol, and is the
gate for any pipeline change: ``tests/test_mbar_protocols.py``.

    python tools/mbar_model.py            # check every protocol listed in PROTOCOLS
"""
from __future__ import annotations

import sys
from collections import deque

# ---------------------------------------------------------------------------------------------
# Model.  A barrier = (completed phases, arrivals in the current phase); `expected` arrivals complete
# a phase.  try_wait.parity(P) is true iff the current (incomplete) phase has parity != P.
# Programs are tuples of ops; an op is a tuple:
#   ("wait", bar, parity)             block until try_wait.parity(parity) holds
#   ("arrive", bar)                   synchronous arrival
#   ("async_arrive", bar)             arrival that happens some time later (tcgen05.commit)
#   ("load", bar, buf, token)         async fill of `buf` with `token`; on landing arrives on `bar`
#   ("read", buf, token)              the agent starts using `buf` and expects `token` in it
#   ("release", buf)                  the agent is done with `buf`
#   ("store", buf, token)             async write of `token` into `buf`, no barrier (a flag-in-payload
#                                     store to a peer: lands whenever)
#   ("poll", buf, token)              block until `buf` holds `token`; a LATER token there means the
#                                     awaited one was overwritten before it was seen (``lost``)
#   ("publish", buf, epoch)           async store of a monotonic epoch (same-address stores of one
#                                     thread stay ordered: a landing never lowers the cell)
#   ("poll_ge", buf, epoch)           block until the cell holds at least `epoch`
#   ("set", buf, token)               synchronous write (local memory, or a peer store ordered before
#                                     the next publish by a system fence + release)
#   ("expect", buf, token)            synchronous read that must find `token` (``stale`` otherwise)
# ---------------------------------------------------------------------------------------------


class Violation(Exception):
    pass


def check(programs: dict, expected: dict, max_states: int = 2_000_000) -> dict:
    """Explore all interleavings.  ``programs``: name -> tuple of ops; ``expected``: barrier ->
    arrival count.  Returns {"states": n} or raises Violation with a trace."""
    names = sorted(programs)
    bars = sorted(expected)
    bidx = {b: i for i, b in enumerate(bars)}
    bufs = sorted({op[2] for p in programs.values() for op in p if op[0] == "load"} |
                  {op[1] for p in programs.values() for op in p
                   if op[0] in ("read", "release", "store", "poll", "publish", "poll_ge", "set", "expect")})
    fidx = {b: i for i, b in enumerate(bufs)}

    # state: (pcs, bar_state, buf_tokens, buf_readers, pending) - all tuples, hashable
    init = (tuple(0 for _ in names),
            tuple((0, 0) for _ in bars),
            tuple(None for _ in bufs),
            tuple(0 for _ in bufs),
            frozenset())
    seen = {init: None}
    queue = deque([init])

    def trace(state, last):
        steps = [last]
        while seen[state] is not None:
            state, step = seen[state]
            steps.append(step)
        return list(reversed(steps))

    def arrive(bar_state, b, state, what):
        phases, got = bar_state[bidx[b]]
        got += 1
        if got > expected[b]:
            raise Violation(("over-arrival", b, trace(state, what)))
        if got == expected[b]:
            phases, got = phases + 1, 0
        bs = list(bar_state)
        bs[bidx[b]] = (phases, got)
        return tuple(bs)

    while queue:
        state = queue.popleft()
        pcs, bar_state, tokens, readers, pending = state
        moves = []
        # asynchronous completions, any order
        for ev in pending:
            moves.append(("async", ev))
        for ai, name in enumerate(names):
            prog = programs[name]
            if pcs[ai] >= len(prog):
                continue
            op = prog[pcs[ai]]
            if op[0] == "wait":
                phases, _ = bar_state[bidx[op[1]]]
                if (phases & 1) == op[2]:
                    continue                          # blocked
            if op[0] == "poll":
                have = tokens[fidx[op[1]]]
                if have is not None and have > op[2]:
                    raise Violation(("lost", op[1], f"holds {have}, still waiting for {op[2]}",
                                     trace(state, (name, op))))
                if have != op[2]:
                    continue                          # blocked
            if op[0] == "poll_ge":
                have = tokens[fidx[op[1]]]
                if have is None or have < op[2]:
                    continue                          # blocked
            moves.append(("step", ai))
        if not moves:
            if any(pcs[ai] < len(programs[n]) for ai, n in enumerate(names)):
                stuck = {n: programs[n][pcs[ai]] for ai, n in enumerate(names) if pcs[ai] < len(programs[n])}
                raise Violation(("deadlock", stuck, trace(state, "stuck")))
            continue
        for mv in moves:
            npcs, nbars, ntok, nrd, npend = list(pcs), bar_state, list(tokens), list(readers), set(pending)
            if mv[0] == "async":
                ev = mv[1]
                npend.discard(ev)
                if ev[0] == "epoch":
                    _, buf, token, _uid = ev
                    have = ntok[fidx[buf]]
                    ntok[fidx[buf]] = token if have is None else max(have, token)
                elif ev[0] == "land":
                    _, b, buf, token, _uid = ev
                    if nrd[fidx[buf]]:
                        raise Violation(("overwrite", buf, trace(state, ev)))
                    ntok[fidx[buf]] = token
                    if b is not None:
                        nbars = arrive(nbars, b, state, ev)
                else:
                    nbars = arrive(nbars, ev[1], state, ev)
                what = ev
            else:
                ai = mv[1]
                name = names[ai]
                op = programs[name][pcs[ai]]
                what = (name, op)
                npcs[ai] += 1
                if op[0] == "arrive":
                    nbars = arrive(nbars, op[1], state, what)
                elif op[0] == "async_arrive":
                    npend.add(("commit", op[1], (name, pcs[ai])))
                elif op[0] == "load":
                    npend.add(("land", op[1], op[2], op[3], (name, pcs[ai])))
                elif op[0] == "store":
                    npend.add(("land", None, op[1], op[2], (name, pcs[ai])))
                elif op[0] == "publish":
                    npend.add(("epoch", op[1], op[2], (name, pcs[ai])))
                elif op[0] == "set":
                    ntok[fidx[op[1]]] = op[2]
                elif op[0] == "expect":
                    if ntok[fidx[op[1]]] != op[2]:
                        raise Violation(("stale", op[1], f"holds {ntok[fidx[op[1]]]}, wanted {op[2]}",
                                         trace(state, what)))
                elif op[0] == "read":
                    if ntok[fidx[op[1]]] != op[2]:
                        raise Violation(("early", op[1], f"holds {ntok[fidx[op[1]]]}, wanted {op[2]}",
                                         trace(state, what)))
                    nrd[fidx[op[1]]] += 1
                elif op[0] == "release":
                    nrd[fidx[op[1]]] -= 1
            nstate = (tuple(npcs), nbars, tuple(ntok), tuple(nrd), frozenset(npend))
            if nstate not in seen:
                seen[nstate] = (state, what)
                if len(seen) > max_states:
                    raise RuntimeError(f"more than {max_states} states: shorten the programs")
                queue.append(nstate)
    return {"states": len(seen)}


# ---------------------------------------------------------------------------------------------
# Protocols (each returns programs, expected)
# ---------------------------------------------------------------------------------------------
def convbn_ring(ring: int = 3, chunks: int = 8, two_full_barriers: bool = True):
    """Shortcut ring of the folded conv+BN GEMM (convbn_tcgen05.cu): one producer fills slots in
    sequence order, the two column halves of epilogue warps take alternate chunks.  With
    ``two_full_barriers=False`` this is the protocol as first written (one "full" barrier per slot)."""
    prod, cons = [], {0: [], 1: []}
    expected = {}
    for s in range(ring):
        expected[f"rempty{s}"] = 1               # one agent stands for the four warps of a half
        if two_full_barriers:
            expected[f"rfull{s}a"] = expected[f"rfull{s}b"] = 1
        else:
            expected[f"rfull{s}"] = 1
    for seq in range(chunks):
        slot, use = seq % ring, seq // ring
        if two_full_barriers:
            full, parity = f"rfull{slot}{'ab'[use & 1]}", (use >> 1) & 1
        else:
            full, parity = f"rfull{slot}", use & 1
        prod += [("wait", f"rempty{slot}", (use & 1) ^ 1), ("load", full, f"slot{slot}", seq)]
        c = cons[seq % 2]
        c += [("wait", full, parity), ("read", f"slot{slot}", seq), ("release", f"slot{slot}"),
              ("arrive", f"rempty{slot}")]
    return {"producer": tuple(prod), "half0": tuple(cons[0]), "half1": tuple(cons[1])}, expected


def operand_pipeline(stages: int = 2, k_blocks: int = 5):
    """TMA producer -> MMA issuer -> (commit) -> producer: the operand ring of every GEMM here."""
    prod, mma = [], []
    expected = {}
    for s in range(stages):
        expected[f"full{s}"] = expected[f"empty{s}"] = 1
    for kb in range(k_blocks):
        s, use = kb % stages, kb // stages
        prod += [("wait", f"empty{s}", (use & 1) ^ 1), ("load", f"full{s}", f"stage{s}", kb)]
        mma += [("wait", f"full{s}", use & 1), ("read", f"stage{s}", kb), ("release", f"stage{s}"),
                ("async_arrive", f"empty{s}")]
    return {"producer": tuple(prod), "mma": tuple(mma)}, expected


def conv3x3_operands(a_stages: int = 2, b_stages: int = 3, strips: int = 2, taps: int = 3,
                     second_issuer_idle: bool = False):
    """conv3x3_tcgen05.cu: a strip producer and a weight producer, two MMA issuers (one per sub-tile)
    that both consume every strip and every weight tile and both release them (count-2 "empty"
    barriers, by tcgen05.commit - or by a plain arrival when the sub-tile has nothing to compute)."""
    pa, pb, iss = [], [], {0: [], 1: []}
    expected = {}
    for s in range(a_stages):
        expected[f"afull{s}"], expected[f"aempty{s}"] = 1, 2
    for s in range(b_stages):
        expected[f"bfull{s}"], expected[f"bempty{s}"] = 1, 2
    bseq = 0
    for cb in range(strips):
        sa, ua = cb % a_stages, cb // a_stages
        pa += [("wait", f"aempty{sa}", (ua & 1) ^ 1), ("load", f"afull{sa}", f"strip{sa}", ("a", cb))]
        for i in (0, 1):
            iss[i] += [("wait", f"afull{sa}", ua & 1), ("read", f"strip{sa}", ("a", cb))]
        for _tap in range(taps):
            sb, ub = bseq % b_stages, bseq // b_stages
            pb += [("wait", f"bempty{sb}", (ub & 1) ^ 1), ("load", f"bfull{sb}", f"w{sb}", ("b", bseq))]
            for i in (0, 1):
                rel = "arrive" if (i == 1 and second_issuer_idle) else "async_arrive"
                iss[i] += [("wait", f"bfull{sb}", ub & 1), ("read", f"w{sb}", ("b", bseq)), ("release", f"w{sb}"),
                           (rel, f"bempty{sb}")]
            bseq += 1
        for i in (0, 1):
            rel = "arrive" if (i == 1 and second_issuer_idle) else "async_arrive"
            iss[i] += [("release", f"strip{sa}"), (rel, f"aempty{sa}")]
    return {"strips": tuple(pa), "weights": tuple(pb), "issuer0": tuple(iss[0]), "issuer1": tuple(iss[1])}, expected


def accumulator_handshake(tiles: int = 5, epilogue_warps: int = 2):
    """MMA issuer <-> epilogue warps over two TMEM accumulator buffers (tfull by commit, tempty by one
    arrival per epilogue warp).  The "buffer" is filled when the commit lands."""
    mma, epi = [], {w: [] for w in range(epilogue_warps)}
    expected = {}
    for a in (0, 1):
        expected[f"tfull{a}"], expected[f"tempty{a}"] = 1, epilogue_warps
    for t in range(tiles):
        a, use = t & 1, t >> 1
        mma += [("wait", f"tempty{a}", (use & 1) ^ 1), ("load", f"tfull{a}", f"acc{a}", t)]
        for w in range(epilogue_warps):
            epi[w] += [("wait", f"tfull{a}", use & 1), ("read", f"acc{a}", t), ("release", f"acc{a}"),
                       ("arrive", f"tempty{a}")]
    progs = {"mma": tuple(mma)}
    progs.update({f"epi{w}": tuple(p) for w, p in epi.items()})
    return progs, expected


def stem_row_ring(stages: int = 5, images: int = 2, oh: int = 4):
    """ResNet stem forward (gemm_tcgen05.cu, conv == 4): input rows stream through the A stages as a
    ring.  Output row `o` of an image needs input rows o .. o + 3, of which only o + 3 is new unless
    the tile starts an image; the issuer waits on all four (three of them for a phase it has already
    seen), releases row o after the tile, and the three rows nobody else will read at the end of an
    image."""
    prod, mma = [], []
    expected = {}
    for s in range(stages):
        expected[f"full{s}"] = expected[f"empty{s}"] = 1
    seq = 0
    base = loaded = 0
    for b in range(images):
        for o in range(oh):
            fresh = o == 0
            for h in range(o if fresh else o + 3, o + 4):
                slot, use = seq % stages, seq // stages
                prod += [("wait", f"empty{slot}", (use & 1) ^ 1), ("load", f"full{slot}", f"row{slot}", (b, h))]
                seq += 1
            base = loaded if fresh else base + 1
            first_new = 0 if fresh else 3
            loaded += 4 if fresh else 1
            for i in range(4):
                sq = base + i
                slot = sq % stages
                mma.append(("wait", f"full{slot}", (sq // stages) & 1))
                mma.append(("read" if i >= first_new else "expect", f"row{slot}", (b, o + i)))
            last = o == oh - 1
            for i in range(4 if last else 1):
                slot = (base + i) % stages
                mma += [("release", f"row{slot}"), ("async_arrive", f"empty{slot}")]
    return {"producer": tuple(prod), "mma": tuple(mma)}, expected


def ll_allreduce(ranks: int = 2, calls: int = 4, parities: int = 2):
    """The flag-in-payload ("LL") all-reduce of comm.cu between kernels of successive calls: every
    rank stores its chunk, tagged with the call number, into its cell of every peer's receive buffer
    of parity call % parities, then polls its own cells for that tag.  Stores land whenever; a
    rank's calls are stream-ordered.  Two parities are enough (a rank cannot start call e + 2 before
    every peer has finished call e); one is not."""
    progs = {}
    for r in range(ranks):
        p = []
        for e in range(calls):
            for q in range(ranks):
                if q != r:
                    p.append(("store", f"rx{q}_from{r}_p{e % parities}", e))
            for q in range(ranks):
                if q != r:
                    p.append(("poll", f"rx{r}_from{q}_p{e % parities}", e))
        progs[f"rank{r}"] = tuple(p)
    return progs, {}


def fused_allreduce(ranks: int = 2, steps: int = 3, closing_barrier: bool = True):
    """The barrier-based fused all-reduce + optimizer kernel of comm.cu (two-shot / NVLS) over
    successive training steps: backward writes the local gradients; barrier 1 (epoch 2s + 1); every
    rank reads its shard of EVERY rank's gradients and stores the new parameters of that shard into
    every rank (ordered before its next publish); barrier 2 (epoch 2s + 2); the rank zeroes its
    gradients; the next forward reads the parameters.  Without the closing barrier a fast rank zeroes
    (and rewrites) gradients a slow peer has not read yet."""
    progs = {}
    for r in range(ranks):
        p = []
        for s in range(steps):
            p.append(("set", f"grad{r}", s))
            for q in range(ranks):
                if q != r:
                    p.append(("publish", f"pad{q}_from{r}", 2 * s + 1))
            for q in range(ranks):
                if q != r:
                    p.append(("poll_ge", f"pad{r}_from{q}", 2 * s + 1))
            for q in range(ranks):
                p.append(("expect", f"grad{q}", s))
            for q in range(ranks):
                p.append(("set", f"param{q}_shard{r}", s))
            if closing_barrier:
                for q in range(ranks):
                    if q != r:
                        p.append(("publish", f"pad{q}_from{r}", 2 * s + 2))
                for q in range(ranks):
                    if q != r:
                        p.append(("poll_ge", f"pad{r}_from{q}", 2 * s + 2))
            p.append(("set", f"grad{r}", "zeroed"))
            for q in range(ranks):
                p.append(("expect", f"param{r}_shard{q}", s))
        progs[f"rank{r}"] = tuple(p)
    return progs, {}


PROTOCOLS = {
    "convbn shortcut ring, 3 slots, two full barriers per slot (as built)": lambda: convbn_ring(3, 8, True),
    "convbn shortcut ring, 4 slots": lambda: convbn_ring(4, 9, True),
    "convbn shortcut ring, 1 slot": lambda: convbn_ring(1, 5, True),
    "GEMM operand ring, 2 stages": lambda: operand_pipeline(2, 6),
    "GEMM operand ring, 3 stages": lambda: operand_pipeline(3, 7),
    "conv3x3 strips + weights, two issuers": lambda: conv3x3_operands(2, 3, 2, 3),
    "conv3x3, second sub-tile idle (plain arrivals)": lambda: conv3x3_operands(2, 2, 2, 2, True),
    "accumulator hand-over, two buffers": lambda: accumulator_handshake(5, 2),
    "stem input-row ring, 5 stages": lambda: stem_row_ring(5, 2, 4),
    "stem input-row ring, 4 stages (the minimum)": lambda: stem_row_ring(4, 2, 3),
    "LL all-reduce, 2 ranks, two receive parities": lambda: ll_allreduce(2, 5, 2),
    "LL all-reduce, 3 ranks, two receive parities": lambda: ll_allreduce(3, 3, 2),
    "fused all-reduce + optimizer, barrier before and after, 2 ranks": lambda: fused_allreduce(2, 3),
    "fused all-reduce + optimizer, 3 ranks, one step": lambda: fused_allreduce(3, 1),
}


def main() -> int:
    bad = 0
    for name, make in PROTOCOLS.items():
        programs, expected = make()
        try:
            res = check(programs, expected)
            print(f"ok    {name}: {res['states']} states")
        except Violation as v:
            bad += 1
            print(f"FAIL  {name}: {v.args[0][0]} {v.args[0][1]}")
    programs, expected = convbn_ring(3, 8, two_full_barriers=False)
    try:
        check(programs, expected)
        print("the one-barrier-per-slot ring passed: the model lost its teeth")
        bad += 1
    except Violation as v:
        kind, where = v.args[0][0], v.args[0][1]
        print(f"ok    convbn shortcut ring as first written (one full barrier per slot) is caught: {kind} on {where}")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
