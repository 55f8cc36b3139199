"""Exhaustive interleaving checks of the mbarrier pipelines of the tcgen05 kernels
(tools/mbar_model.py): every protocol as built passes; the shortcut ring as first written - the race
that crashed the folded conv+BN GEMM once in a few thousand launches - is
found from the protocol alone.  (the reference has no race detection of its
own; compute-sanitizer is closed on the GPU pool and does not model phase parity.)"""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import mbar_model as mm  # noqa: E402


@pytest.mark.parametrize("name", sorted(mm.PROTOCOLS))
def test_protocol_as_built_has_no_bad_interleaving(name):
    programs, expected = mm.PROTOCOLS[name]()
    assert mm.check(programs, expected)["states"] > 10


@pytest.mark.parametrize("ring", [1, 3, 5])
def test_one_full_barrier_per_slot_is_unsound_for_odd_rings(ring):
    # fills of a slot alternate between the two column halves: a half sees every other phase of the
    # slot's barrier and takes "two phases ago" for "done"
    programs, expected = mm.convbn_ring(ring, 2 * ring + 4, two_full_barriers=False)
    with pytest.raises(mm.Violation) as e:
        mm.check(programs, expected)
    assert e.value.args[0][0] in ("early", "over-arrival", "overwrite", "deadlock")


@pytest.mark.parametrize("ring", [2, 4])
def test_one_full_barrier_per_slot_is_sound_for_even_rings(ring):
    # an even ring gives each half its own slots: every waiter sees every phase
    programs, expected = mm.convbn_ring(ring, 2 * ring + 3, two_full_barriers=False)
    mm.check(programs, expected)


def test_the_checker_sees_a_missing_release_and_a_double_arrival():
    programs, expected = mm.operand_pipeline(2, 4)
    broken = dict(programs)
    broken["mma"] = tuple(op for op in programs["mma"] if op[0] != "async_arrive")      # never frees a stage
    with pytest.raises(mm.Violation) as e:
        mm.check(broken, expected)
    assert e.value.args[0][0] == "deadlock"
    twice = dict(programs)
    mma = list(programs["mma"])
    mma.insert(4, ("arrive", "empty0"))
    twice["mma"] = tuple(mma)
    with pytest.raises(mm.Violation):
        mm.check(twice, expected)


def test_a_consumer_that_skips_the_wait_is_caught_as_an_early_read():
    programs, expected = mm.accumulator_handshake(3, 2)
    epi = [op for op in programs["epi1"]]
    del epi[0]                                             # reads the first accumulator without waiting
    broken = dict(programs)
    broken["epi1"] = tuple(epi)
    with pytest.raises(mm.Violation) as e:
        mm.check(broken, expected)
    assert e.value.args[0][0] == "early"


def test_ll_allreduce_needs_two_receive_parities():
    # flag-in-payload all-reduce of comm.cu: with one receive buffer a fast rank's store of call e + 1
    # overwrites the cell a slow peer is still polling for call e
    with pytest.raises(mm.Violation) as e:
        mm.check(*mm.ll_allreduce(2, 4, parities=1))
    assert e.value.args[0][0] == "lost"
    mm.check(*mm.ll_allreduce(2, 6, parities=2))


def test_fused_allreduce_needs_its_closing_barrier():
    # comm.cu, barrier-based schedules: without the second barrier a rank reads parameters a peer has
    # not stored yet (and would zero gradients a peer has not read yet)
    with pytest.raises(mm.Violation) as e:
        mm.check(*mm.fused_allreduce(2, 2, closing_barrier=False))
    assert e.value.args[0][0] == "stale"
    mm.check(*mm.fused_allreduce(2, 3, closing_barrier=True))


def test_stem_row_ring_needs_four_stages():
    # a tile needs four input rows at once: with three stages the producer and the issuer wait on
    # each other for ever
    with pytest.raises(mm.Violation) as e:
        mm.check(*mm.stem_row_ring(3, 1, 3))
    assert e.value.args[0][0] == "deadlock"
    mm.check(*mm.stem_row_ring(8, 2, 4))              # the kernel's depth
