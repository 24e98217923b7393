"""Host-side model of the two-edge ATM kernel's visiting schedule (dftd4_b200/csrc/d4_atm.cuh, atm_two_edge_kernel).

The kernel's correctness rests on a counting argument: with a centre's row ordered [owned partners | others], every
unordered pair of owned partners must be taken by exactly ONE lane (which then produces both centre edges), and every
(owned, other) pair exactly once by the owned partner's lane.  The index arithmetic that decides this (pass / chunk /
warp start offsets, triangular or cyclic pairing inside a pass's own 256-entry block) is restated here loop for loop and
checked exhaustively for row shapes around every boundary (pass size 256, chunk size 128, warp size 32).
No GPU and no oracle involved: this is a test of the schedule, not of the arithmetic.
"""
from collections import Counter

import pytest

THREADS, KC, WARP = 256, 128, 32


def schedule(n, n_own, cyclic):
    """Yield ('two', pj, pk) / ('one', pj, pk) for every accepted (lane, third entry) of one centre; positions are row
    positions after the owned-first partition.  Mirrors the loops of atm_two_edge_kernel with all distance tests true."""
    for p0 in range(0, n_own, THREADS):                       # passes: 256 consecutive owned entries
        for c0 in range(p0, n, KC):                           # chunks from the pass's own position on
            kc = min(KC, n - c0)
            k2end = min(kc, n_own - c0)
            for warp in range(THREADS // WARP):
                w0 = p0 + warp * WARP
                lanes = [(lane, w0 + lane) for lane in range(WARP) if w0 + lane < n_own]   # valid lanes
                if c0 - p0 >= THREADS:
                    rng = range(0, max(k2end, 0))
                    for kk in rng:
                        for _, pj in lanes:
                            yield ("two", pj, c0 + kk)
                elif cyclic:
                    half = THREADS // 2
                    for kk in range(0, max(k2end, 0)):
                        t = (c0 + kk - w0) & (THREADS - 1)
                        if t < 1 or t > half + 31:
                            continue
                        for lane, pj in lanes:
                            d = (t - lane) & (THREADS - 1)
                            if d >= 1 and (d < half or (d == half and pj < c0 + kk)):
                                yield ("two", pj, c0 + kk)
                else:
                    for kk in range(max(0, w0 + 1 - c0), max(k2end, 0)):
                        for _, pj in lanes:
                            if c0 + kk > pj:
                                yield ("two", pj, c0 + kk)
                for kk in range(max(0, n_own - c0), kc):      # partners above the centre: one-edge visits
                    for _, pj in lanes:
                        yield ("one", pj, c0 + kk)


SHAPES = [(0, 0), (5, 0), (1, 1), (2, 2), (40, 31), (40, 32), (40, 33), (200, 127), (200, 128), (200, 129),
          (300, 255), (300, 256), (300, 257), (700, 300), (700, 383), (700, 384), (700, 385), (700, 513), (700, 700),
          (1000, 640), (257, 257), (513, 512)]


@pytest.mark.parametrize("cyclic", [False, True], ids=["triangular", "cyclic"])
@pytest.mark.parametrize("n,n_own", SHAPES)
def test_every_pair_is_visited_exactly_once(n, n_own, cyclic):
    two, one = Counter(), Counter()
    for kind, pj, pk in schedule(n, n_own, cyclic):
        assert pj < n_own and pk < n and pj != pk
        if kind == "two":
            assert pk < n_own
            two[frozenset((pj, pk))] += 1
        else:
            assert pk >= n_own
            one[(pj, pk)] += 1
    assert len(two) == n_own * (n_own - 1) // 2 and all(v == 1 for v in two.values())
    assert len(one) == n_own * (n - n_own) and all(v == 1 for v in one.values())


def test_second_edge_sums_are_complete_before_an_entry_is_finalised():
    """An entry's accumulator is read after the barrier that ends ITS pass: every lane that takes it as second edge must
    belong to the same or an earlier pass (both pairing rules)."""
    for cyclic in (False, True):
        for kind, pj, pk in schedule(900, 700, cyclic):
            if kind == "two":
                assert pj // THREADS <= pk // THREADS
