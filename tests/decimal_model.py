"""Exact-rational model of the Decimal contract restated in oracle/dec96.h.

Independent second opinion on oracle/dec96.c: Python big integers and `fractions`, no shared
code with the C restatement.  value = m / 10**scale with m < 2**96 and 0 <= scale <= 28; every
operator returns the exact result rounded half-to-even at the largest scale whose mantissa
fits (rust_decimal 1.39.0's documented format; the crate source is not under /root/reference,
so this is "parity unpinned" -- see DESIGN.md).
"""
from fractions import Fraction

MAX_MANTISSA = 2**96 - 1
MAX_SCALE = 28


def _round_half_even(fr: Fraction) -> int:
    fl = fr.numerator // fr.denominator
    rem = fr - fl
    if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and fl % 2 == 1):
        return fl + 1
    return fl


def fit(fr: Fraction) -> Fraction:
    """Round an exact rational to the nearest representable decimal (largest scale that fits)."""
    sign = -1 if fr < 0 else 1
    fr = abs(fr)
    for scale in range(MAX_SCALE, -1, -1):
        m = _round_half_even(fr * 10**scale)
        if m <= MAX_MANTISSA:
            return sign * Fraction(m, 10**scale)
    raise OverflowError("decimal overflow")


def parse(s: str) -> Fraction:
    return fit(Fraction(s))


def to_str(fr: Fraction) -> str:
    """Value-canonical positional string (what dec96_to_str_normalized prints)."""
    sign = "-" if fr < 0 else ""
    fr = abs(fr)
    for scale in range(0, MAX_SCALE + 1):
        if (fr * 10**scale).denominator == 1:
            m = int(fr * 10**scale)
            digits = str(m).rjust(scale + 1, "0")
            if scale == 0:
                return sign + digits
            return sign + digits[:-scale] + "." + digits[-scale:]
    raise ValueError("not a decimal")


def add(a, b):
    return fit(a + b)


def sub(a, b):
    return fit(a - b)


def mul(a, b):
    return fit(a * b)


def div(a, b):
    return fit(a / b)


def tally(n, votes, weights):
    """client.rs:385-449 with `votes[m]` a list of Fractions or None and `weights[m]` a Fraction
    or None.  Returns (choice_weight, confidence, voter_confidence-or-None)."""
    cw = [Fraction(0)] * n
    for vote, weight in zip(votes, weights):
        if vote is None:
            continue
        wm = weight if weight is not None else Fraction(0)
        for i, v in enumerate(vote):
            cw[i] = add(cw[i], mul(v, wm))
    total = Fraction(0)
    for c in cw:
        total = add(total, c)
    conf = [div(c, total) if total > 0 else Fraction(0) for c in cw]
    vconf = []
    for vote in votes:
        if vote is None:
            vconf.append(None)
            continue
        acc = Fraction(0)
        for i, v in enumerate(vote):
            q = div(cw[i], total) if total > 0 else Fraction(0)
            acc = add(acc, mul(q, v))
        vconf.append(acc)
    return cw, conf, vconf


def merge_error_codes(codes):
    """client.rs:386-409,459-463 + score/completions/error.rs:41. 0 = voter without error."""
    all_error, code = True, None
    for e in codes:
        if not all_error:
            break
        if e == 0:
            all_error = False
        elif code is None:
            code = e
        elif e != code:
            code = 400 if (400 <= e < 500 and 400 <= code < 500) else 500
    if not all_error:
        return 0
    return code if code is not None else 500
