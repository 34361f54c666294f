"""oracle/dec96.c against the exact-rational model (tests/decimal_model.py).

The reference's arithmetic lives in rust_decimal 1.39.0 (Cargo.lock:1310-1311), absent from
/root/reference: PARITY UNPINNED.  These tests pin the C restatement to the published number
format (96-bit mantissa, scale <= 28, half-even) -- they cannot pin it to the crate itself.
"""
import random
from fractions import Fraction

import pytest

import decimal_model as dm


def rand_decimal(rng, big=False):
    digits = rng.choice([1, 2, 3, 5, 9, 12, 18, 24, 28, 29] if big else [1, 2, 3, 4, 6, 9])
    m = rng.randrange(0, 10**digits)
    m = min(m, dm.MAX_MANTISSA)
    scale = rng.randrange(0, 29)
    sign = rng.choice([1, 1, 1, -1])
    return sign * Fraction(m, 10**scale)


def test_known_answers(oracle):
    assert oracle.dec_op("1", "3", "/") == "0.3333333333333333333333333333"
    assert oracle.dec_op("2", "3", "/") == "0.6666666666666666666666666667"
    # 29 significant digits fit when the leading digits stay under 7.92...e28
    assert oracle.dec_op("5", "3", "/") == "1.6666666666666666666666666667"
    assert oracle.dec_op("25", "3", "/") == "8.333333333333333333333333333"
    assert oracle.dec_op("1", "2", "/") == "0.5"
    assert oracle.dec_op("0.1", "0.2", "+") == "0.3"
    assert oracle.dec_op("1.10", "2", "*") == "2.2"
    assert oracle.dec_op("79228162514264337593543950335", "0", "+") == "79228162514264337593543950335"
    assert oracle.dec_op("0.0000000000000000000000000001", "0.1", "*") == "0"  # half-even at scale 28
    assert oracle.dec_op("0.0000000000000000000000000003", "0.5", "*") == "0.0000000000000000000000000002"
    assert oracle.dec_op("0.0000000000000000000000000005", "0.5", "*") == "0.0000000000000000000000000002"
    assert oracle.dec_op("0.0000000000000000000000000007", "0.5", "*") == "0.0000000000000000000000000004"
    assert oracle.dec_op("2", "2.0", "cmp") == "0"
    assert oracle.dec_op("-1", "0.5", "cmp") == "-1"
    with pytest.raises(ZeroDivisionError):
        oracle.dec_op("1", "0", "/")
    with pytest.raises(OverflowError):
        oracle.dec_op("79228162514264337593543950335", "1", "+")
    with pytest.raises(OverflowError):
        oracle.dec_op("79228162514264337593543950335", "1.1", "*")


@pytest.mark.parametrize("op", ["+", "-", "*", "/"])
@pytest.mark.parametrize("big", [False, True])
def test_random_against_exact_rational(oracle, op, big):
    rng = random.Random(hash((op, big)) & 0xFFFF)
    fn = {"+": dm.add, "-": dm.sub, "*": dm.mul, "/": dm.div}[op]
    checked = 0
    for _ in range(1500):
        a, b = rand_decimal(rng, big), rand_decimal(rng, big)
        if op == "/" and b == 0:
            continue
        try:
            want = dm.to_str(fn(a, b))
        except OverflowError:
            with pytest.raises(OverflowError):
                oracle.dec_op(dm.to_str(a), dm.to_str(b), op)
            continue
        got = oracle.dec_op(dm.to_str(a), dm.to_str(b), op)
        assert got == want, (dm.to_str(a), op, dm.to_str(b))
        checked += 1
    assert checked > 1000


def test_parse_rounds_long_literals(oracle):
    # more than 28 fractional digits: half-even at the last representable place
    assert oracle.dec_op("0.33333333333333333333333333335", "0", "+") == "0.3333333333333333333333333334"
    assert oracle.dec_op("0.33333333333333333333333333325", "0", "+") == "0.3333333333333333333333333332"
    assert oracle.dec_op("1e-7", "0", "+") == "0.0000001"
    assert oracle.dec_op("2.5E+3", "0", "+") == "2500"


def test_float_bridge(oracle):
    assert oracle.dec_from_f32(0.1) == "0.1"
    assert oracle.dec_from_f32(1.5) == "1.5"
    assert oracle.dec_from_f32(0.51) == "0.51"
    assert oracle.dec_to_double("0.6666666666666666666666666667") == 2.0 / 3.0
    assert oracle.dec_to_double("0.1") == 0.1
