"""The oracle's restatement of client.rs:384-463 against the committed known answers
(tests/golden/tally_kats.json, derived by tests/golden/make_golden.py -- the reference has no
golden vectors of its own: PARITY UNPINNED) and the internal consistency of its f64 stages."""
import json
import os

import numpy as np
import pytest

import decimal_model as dm
from parity_util import check_rank, check_winner

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "tally_kats.json")))


def to_arrays(case, n, with_E=False):
    votes, weights = case["votes"], case["weights"]
    M = len(votes)
    V = np.zeros((1, M, n), np.float32)
    w = np.zeros((1, M), np.float32)
    for m in range(M):
        if votes[m] is None:
            V[0, m, :] = np.nan
        else:
            V[0, m] = [float(x) for x in votes[m]]
        w[0, m] = np.nan if weights[m] is None else float(weights[m])
    return V, w


@pytest.mark.parametrize("case", GOLDEN["tally"], ids=lambda c: c["name"])
def test_decimal_tally_matches_golden(oracle, case):
    cw, conf, vconf = oracle.tally_decimal(case["n_choices"], case["votes"], case["weights"])
    assert cw == case["choice_weight"]
    assert conf == case["confidence"]
    assert vconf == case["voter_confidence"]
    if "error_codes" in case:
        assert oracle.merge_error_codes(case["error_codes"]) == case["merged_status"]


@pytest.mark.parametrize("case", GOLDEN["tally"], ids=lambda c: c["name"])
def test_f64_identity_mode_is_the_reference_tally(oracle, case):
    """similarity = IDENTITY must reproduce the Decimal tally (the S = I bridge)."""
    n = case["n_choices"]
    V, w = to_arrays(case, n)
    for fn in (oracle.score_f64, oracle.score_port):
        out = fn(np.zeros((1, n, 0), np.float32), V, w, oracle.SIM_IDENTITY)
        want_cw = np.array([float(x) for x in case["choice_weight"]])
        want_cf = np.array([float(x) for x in case["confidence"]])
        np.testing.assert_allclose(out["cw"][0], want_cw, rtol=1e-6, atol=1e-12)
        np.testing.assert_allclose(out["conf"][0], want_cf, rtol=1e-6, atol=1e-12)
        # inputs pass through fp32 here, so 28-digit near-ties are not resolvable
        check_winner(want_cf, case["winner"], out["winner"][0])
        check_rank(want_cf, case["rank"], out["rank"][0])
        for m, vc in enumerate(case["voter_confidence"]):
            if vc is None:
                assert np.isnan(out["agree"][0, m])
            else:
                assert out["agree"][0, m] == pytest.approx(float(vc), rel=1e-6, abs=1e-12)


@pytest.mark.parametrize("case", GOLDEN["cosine"], ids=lambda c: c["name"])
def test_f64_cosine_matches_golden(oracle, case):
    E = np.array([[float(x) for x in row] for row in case["E"]], np.float32)[None]
    n = E.shape[1]
    V, w = to_arrays(case, n)
    for fn in (oracle.score_f64, oracle.score_port):
        out = fn(E, V, w, oracle.SIM_COSINE)
        # the oracle takes the fp32 inputs the device sees: 5e-7 covers rounding 0.7 etc. to fp32
        np.testing.assert_allclose(out["cw"][0], case["choice_weight"], rtol=5e-7, atol=1e-9)
        np.testing.assert_allclose(out["cw"][0], case["choice_weight_centroid"], rtol=5e-7, atol=1e-9)
        np.testing.assert_allclose(out["conf"][0], case["confidence"], rtol=5e-7, atol=1e-9)
        assert out["winner"][0] == case["winner"]
        assert list(out["rank"][0]) == case["rank"]
        for m, vc in enumerate(case["voter_confidence"]):
            if vc is None:
                assert np.isnan(out["agree"][0, m])
            else:
                assert out["agree"][0, m] == pytest.approx(vc, rel=5e-7)


def test_merge_error_codes_against_model(oracle):
    rng = np.random.default_rng(7)
    for _ in range(500):
        codes = [int(x) for x in rng.choice([0, 400, 404, 429, 500, 503], size=rng.integers(0, 6))]
        assert oracle.merge_error_codes(codes) == dm.merge_error_codes(codes)


def test_random_tally_against_model(oracle):
    from fractions import Fraction

    rng = np.random.default_rng(11)
    for _ in range(60):
        n, M = int(rng.integers(2, 9)), int(rng.integers(1, 6))
        votes, weights = [], []
        for m in range(M):
            if rng.random() < 0.15:
                votes.append(None)
            else:
                raw = rng.integers(0, 1000, size=n)
                raw[rng.integers(0, n)] += 1
                tot = int(raw.sum())
                votes.append([dm.to_str(dm.div(Fraction(int(x)), Fraction(tot))) for x in raw])
            weights.append(None if rng.random() < 0.1 else "%d.%02d" % (rng.integers(0, 3), rng.integers(0, 100)))
        fv = [None if v is None else [dm.parse(x) for x in v] for v in votes]
        fw = [None if x is None else dm.parse(x) for x in weights]
        cw, conf, vconf = dm.tally(n, fv, fw)
        got = oracle.tally_decimal(n, votes, weights)
        assert got[0] == [dm.to_str(x) for x in cw]
        assert got[1] == [dm.to_str(x) for x in conf]
        assert got[2] == [None if x is None else dm.to_str(x) for x in vconf]


def test_pool_centroid_identity(oracle):
    """K9 at scale: explicit pairwise matrix == centroid form to 1e-12 (f64)."""
    rng = np.random.default_rng(3)
    N, D, M = 96, 40, 5
    E = rng.standard_normal((N, D)).astype(np.float32)
    E[7] = 0
    V = rng.random((M, N)).astype(np.float32)
    V[2] = np.nan
    w = rng.uniform(0.5, 2.0, M).astype(np.float32)
    a = oracle.pool_f64(E, V, w, method=0)
    b = oracle.pool_f64(E, V, w, method=1)
    np.testing.assert_allclose(a["cw"], b["cw"], rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(a["conf"], b["conf"], rtol=1e-11, atol=1e-13)
    assert a["winner"] == b["winner"]
    # and the pool form agrees with the per-request form on the same data
    c = oracle.score_f64(E[None], V[None], w[None], oracle.SIM_COSINE)
    np.testing.assert_allclose(a["cw"], c["cw"][0], rtol=1e-11, atol=1e-12)
    assert a["winner"] == c["winner"][0]


def test_vote_from_logprobs(oracle):
    # client.rs:1765-1791: three alternatives on sibling leaves 2,0,2 and one foreign token
    lp = [-0.1, -2.5, -3.0, -0.7]
    leaf = [2, 0, 2, -1]
    v = oracle.vote_from_logprobs(4, lp, leaf)
    p = np.exp(np.array(lp[:3]))
    want = np.array([p[1], 0, p[0] + p[2], 0]) / p.sum()
    np.testing.assert_allclose(v, want, rtol=1e-15)
    assert v.sum() == pytest.approx(1.0, abs=1e-15)
