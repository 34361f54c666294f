/*
 * dec96 -- TEST INFRASTRUCTURE ONLY (CPU oracle).  Not part of the product path.
 *
 * A small restatement of the arithmetic the reference's tally executes inside the
 * third-party crate rust_decimal 1.39.0 (Cargo.lock:1310-1311; features serde-float,
 * macros, maths, Cargo.toml:28).  That crate's source is NOT under /root/reference, so
 * what is restated here is its published number format and rounding contract:
 *
 *   value = (-1)^neg * m / 10^scale,  m an unsigned 96-bit integer, 0 <= scale <= 28.
 *   +, -, *  : exact when the exact result is representable; otherwise the exact result
 *              is rounded half-to-even to the largest scale (<= 28) whose mantissa still
 *              fits 96 bits.
 *   /        : the quotient is carried to the largest scale (<= 28) whose mantissa fits
 *              96 bits and rounded half-to-even on the exact remainder.
 *
 * PARITY UNPINNED: the reference ships no tests or golden vectors for this path
 *  and cannot be compiled here (no Rust toolchain), so these semantics
 * are pinned only against an independent exact-rational model (tests/test_oracle_dec96.py)
 * and hand-derived known answers (tests/golden/).
 *
 * Call sites in the reference that these operators serve:
 *   Decimal::ZERO init ........ src/score/completions/client.rs:385
 *   *  and +=  ................ src/score/completions/client.rs:412-413
 *   .sum::<Decimal>() ......... src/score/completions/client.rs:422-424
 *   > and /  .................. src/score/completions/client.rs:431-435, 440-444
 *   * and +  .................. src/score/completions/client.rs:444-447
 */
#ifndef LWC_ORACLE_DEC96_H
#define LWC_ORACLE_DEC96_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    uint32_t m[3]; /* little-endian limbs of the 96-bit mantissa */
    uint8_t scale; /* 0..28 */
    uint8_t neg;   /* 1 = negative (zero is always stored non-negative) */
} dec96;

enum { DEC96_OK = 0, DEC96_OVERFLOW = 1, DEC96_DIV_BY_ZERO = 2, DEC96_PARSE = 3 };

dec96 dec96_zero(void);
dec96 dec96_one(void);
int dec96_is_zero(const dec96 *a);

/* Parse "[-]digits[.digits]" (more than 28 fractional digits / 96 bits are rounded
 * half-even, as rust_decimal's from_str does). */
int dec96_from_str(const char *s, dec96 *out);
/* Plain positional notation, trailing zeros of the stored scale kept. Returns length. */
size_t dec96_to_str(const dec96 *a, char *buf, size_t buflen);
/* Same but with trailing fractional zeros stripped (value-canonical form). */
size_t dec96_to_str_normalized(const dec96 *a, char *buf, size_t buflen);
/* Nearest double (through the decimal string, correctly rounded by strtod). */
double dec96_to_double(const dec96 *a);
/* Shortest decimal string that round-trips the float/double, parsed as a decimal.
 * This mirrors how a JSON float lands in a Decimal under serde-float. */
int dec96_from_f32(float x, dec96 *out);
int dec96_from_f64(double x, dec96 *out);

int dec96_add(const dec96 *a, const dec96 *b, dec96 *out);
int dec96_sub(const dec96 *a, const dec96 *b, dec96 *out);
int dec96_mul(const dec96 *a, const dec96 *b, dec96 *out);
int dec96_div(const dec96 *a, const dec96 *b, dec96 *out);
/* -1, 0, +1 by numeric value (scale-insensitive). */
int dec96_cmp(const dec96 *a, const dec96 *b);

#ifdef __cplusplus
}
#endif
#endif
