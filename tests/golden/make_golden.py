"""Generates tests/golden/tally_kats.json -- the known-answer tests of SURVEY.md 8c (K1-K11).

The reference holds no golden vectors for this path and cannot be run here (Rust, no
toolchain), so these answers are DERIVED from the cited lines of
src/score/completions/client.rs:384-463 with the exact-rational model in
tests/decimal_model.py (Python big integers; no code shared with oracle/*.c), and for the
build-defined cosine stages with Python `decimal` at 60 digits.  Re-run:
    python tests/golden/make_golden.py
"""
import json
import os
import sys
from decimal import Decimal, getcontext
from fractions import Fraction

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import decimal_model as dm  # noqa: E402

getcontext().prec = 60


def tally_case(name, n, votes, weights, note, codes=None):
    fv = [None if v is None else [dm.parse(x) for x in v] for v in votes]
    fw = [None if w is None else dm.parse(w) for w in weights]
    cw, conf, vconf = dm.tally(n, fv, fw)
    total = sum(cw)
    winner = -1
    if total > 0:
        winner = max(range(n), key=lambda i: (conf[i], -i))
    rank = sorted(range(n), key=lambda i: (-conf[i], i))
    case = {
        "name": name, "note": note, "n_choices": n, "votes": votes, "weights": weights,
        "choice_weight": [dm.to_str(x) for x in cw],
        "confidence": [dm.to_str(x) for x in conf],
        "voter_confidence": [None if x is None else dm.to_str(x) for x in vconf],
        "winner": winner, "rank": rank,
    }
    if codes is not None:
        case["error_codes"] = codes
        case["merged_status"] = dm.merge_error_codes(codes)
    return case


def cosine_case(name, E, votes, weights, note):
    """Build-defined stages at 60-digit precision on exactly-representable inputs."""
    n = len(E)
    Ed = [[Decimal(x) for x in row] for row in E]
    norms = [sum(x * x for x in row).sqrt() for row in Ed]
    t = [Decimal(0)] * n
    for v, w in zip(votes, weights):
        if v is None:
            continue
        wm = Decimal(w) if w is not None else Decimal(0)
        for i, x in enumerate(v):
            t[i] += Decimal(x) * wm
    S = [[(sum(a * b for a, b in zip(Ed[i], Ed[j])) / (norms[i] * norms[j])) if norms[i] and norms[j] else Decimal(0)
          for j in range(n)] for i in range(n)]
    cw = [sum(t[i] * S[i][j] for i in range(n)) for j in range(n)]
    total = sum(cw)
    conf = [c / total if total > 0 else Decimal(0) for c in cw]
    agree = [None if v is None else float(sum(conf[i] * Decimal(x) for i, x in enumerate(v))) for v in votes]
    # centroid identity (K9): cw_j = <E^_j, sum_i t_i E^_i>
    D = len(E[0])
    cvec = [sum(t[i] * Ed[i][k] / norms[i] for i in range(n) if norms[i]) for k in range(D)]
    cw_centroid = [float(sum(Ed[j][k] * cvec[k] for k in range(D)) / norms[j]) if norms[j] else 0.0 for j in range(n)]
    winner = max(range(n), key=lambda i: (conf[i], -i)) if total > 0 else -1
    return {
        "name": name, "note": note, "E": E, "votes": votes, "weights": weights,
        "choice_weight": [float(c) for c in cw], "confidence": [float(c) for c in conf],
        "voter_confidence": agree, "choice_weight_centroid": cw_centroid,
        "winner": winner, "rank": sorted(range(n), key=lambda i: (-conf[i], i)),
    }


def main():
    tally = [
        tally_case("K1_one_hot", 3, [["1", "0", "0"], ["0", "1", "0"]], ["1", "2"],
                   "cw=[1,2,0]; conf=[1/3,2/3,0] at 28 digits; winner 1 (client.rs:410-437)"),
        tally_case("K2_all_failed_429", 3, [None, None], ["1", "2"],
                   "no vote rows: sum=0 -> all confidences 0 (client.rs:431-435)", codes=[429, 429]),
        tally_case("K2_all_failed_400_429", 2, [None, None], ["1", "1"], "4xx mismatch -> 400", codes=[400, 429]),
        tally_case("K2_all_failed_400_500", 2, [None, None], ["1", "1"], "mixed -> 500", codes=[400, 500]),
        tally_case("K2_all_failed_500_404", 2, [None, None], ["1", "1"], "mixed -> 500", codes=[500, 404]),
        tally_case("K2_one_succeeded", 2, [["1", "0"], None], ["1", "1"], "first voter fine -> no error", codes=[0, 503]),
        tally_case("K3_weight_none", 3, [["1", "0", "0"], ["0", "0", "1"]], [None, "2.5"],
                   "weight None contributes 0 (client.rs:412-413)"),
        tally_case("K4_probabilistic", 3, [["0.7", "0.2", "0.1"], ["0.1", "0.1", "0.8"]], ["1.5", "0.5"],
                   "cw=[1.10,0.35,0.55]"),
        tally_case("K5_error_after_vote", 2, [["0.25", "0.75"], ["1", "0"]], ["1", "3"],
                   "vote presence is the only gate (client.rs:410): an errored voter with a vote still counts",
                   codes=[500, 0]),
        tally_case("K10_exact_tie", 4, [["1", "0", "0", "0"], ["0", "0", "1", "0"]], ["1.25", "1.25"],
                   "choices 0 and 2 tie exactly -> lowest index wins; rank is stable"),
        tally_case("K_thirds_28_digits", 3, [["0.3333333333333333333333333333", "0.3333333333333333333333333333",
                                             "0.3333333333333333333333333334"]], ["1.37"],
                   "28-digit votes x 2-digit weight: product scale 30 rounds half-even to 28"),
        tally_case("K_sevenths", 7, [[dm.to_str(dm.div(Fraction(k + 1), Fraction(28))) for k in range(7)],
                                     [dm.to_str(dm.div(Fraction(7 - k), Fraction(28))) for k in range(7)]],
                   ["0.51", "1.99"], "quotient votes as get_vote produces them (client.rs:1789-1791)"),
    ]
    # K11: N=128, <=20 non-zeros per vote row (prefix-tree branch width, client.rs:504-507)
    n = 128
    rows = []
    for m in range(16):
        start = (m * 37) % (n - 20)
        raw = [Fraction((7 * k + 3 * m) % 11 + 1) for k in range(20)]
        tot = sum(raw)
        row = ["0"] * n
        for k in range(20):
            row[start + k] = dm.to_str(dm.div(raw[k], tot))
        rows.append(row)
    rows[5] = None
    tally.append(tally_case("K11_n128_sparse", n, rows, ["%d.%02d" % (1 + (m % 2), (13 * m) % 100) for m in range(16)],
                            "N=128, 16 voters, 20 contiguous non-zeros per row, one voter absent"))

    eye4 = [["1" if i == j else "0" for j in range(4)] for i in range(4)]
    cosine = [
        cosine_case("K6_orthonormal_is_identity", [[x for x in row] for row in eye4][:3],
                    [["1", "0", "0"], ["0", "1", "0"]], ["1", "2"],
                    "orthonormal E makes S = I, so cosine mode reproduces K1 (SURVEY 8.0 bridge)"),
        cosine_case("K6b_orthogonal_scaled", [["3", "0", "0", "0"], ["0", "0.5", "0", "0"], ["0", "0", "0", "7"]],
                    [["0.7", "0.2", "0.1"], ["0.1", "0.1", "0.8"]], ["1.5", "0.5"],
                    "orthogonal rows of different length: normalisation must make S = I (reproduces K4)"),
        cosine_case("K7_duplicate_rows_pool_votes", [["1", "2", "2"], ["1", "2", "2"], ["2", "-2", "1"]],
                    [["1", "0", "0"], ["0", "1", "0"], ["0", "0", "1"]], ["1", "1", "1.5"],
                    "rows 0 and 1 identical (S_01 = 1): both get weight 2, the orthogonal row 2 keeps 1.5"),
        cosine_case("K8_zero_row", [["0", "0", "0"], ["1", "1", "0"], ["1", "0", "1"]],
                    [["0.5", "0.25", "0.25"], ["0", "1", "0"]], ["1", "1"],
                    "zero-norm row 0: E^_0 = 0, its column of S is 0, votes cast on it are lost"),
        cosine_case("K9_general", [["1", "2", "3", "4"], ["4", "3", "2", "1"], ["-1", "0.5", "2", "0.25"],
                                   ["0.125", "8", "-3", "1"], ["2", "2", "2", "2"]],
                    [["0.1", "0.2", "0.3", "0.15", "0.25"], ["0", "0", "1", "0", "0"], None,
                     ["0.5", "0", "0", "0.5", "0"]], ["0.75", "1.25", "2", None],
                    "general case incl. an absent voter and a None weight; also pins the centroid identity"),
    ]
    out = {"comment": "generated by tests/golden/make_golden.py; see its docstring", "tally": tally, "cosine": cosine}
    with open(os.path.join(HERE, "tally_kats.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", len(tally), "tally and", len(cosine), "cosine KATs")


if __name__ == "__main__":
    main()
